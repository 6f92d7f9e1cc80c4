#!/bin/bash
# aggregation kernel variants + the C2-size comparison against the live reference extension.  usage: tools/gpu_pass4.sh <tag>
set -u
TAG=${1:-rXX}
OUT=gpurun_out/$TAG
mkdir -p $OUT
for v in gpurun_variants/*.so; do
  n=$(basename $v .so)
  PIGS_B200_BINDING=ctypes PIGS_B200_LIB=$PWD/$v timeout 300 python tools/bench_hbm.py --only=agg > $OUT/hbm_$n.json 2> $OUT/hbm_$n.err
  python -c "
import json
for r in json.load(open('$OUT/hbm_$n.json'))['rows'][1:]: print('$n', r['kernel'][:50], r['ms'], r['frac_of_hbm_peak'])"
done
PIGS_RUN_SLOW=1 timeout 900 python -m pytest tests/test_gpu_vs_reference.py -m gpu -q -s -k c2_size > $OUT/ref_c2_parity.log 2>&1; echo "ref c2 exit $?"
tail -12 $OUT/ref_c2_parity.log
