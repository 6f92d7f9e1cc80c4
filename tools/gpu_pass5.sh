#!/bin/bash
# aggregation forward grid sweep + ncu capture of the neighbour-search kernels.  usage: tools/gpu_pass5.sh <tag>
set -u
TAG=${1:-rXX}
OUT=gpurun_out/$TAG
mkdir -p $OUT
for w in 1 2 3; do
  PIGS_A16_FWD_WAVES=$w timeout 300 python tools/bench_hbm.py --only=agg > $OUT/hbm_waves$w.json 2> $OUT/hbm_waves$w.err
  python -c "
import json
for r in json.load(open('$OUT/hbm_waves$w.json'))['rows'][1:]: print('waves $w', r['kernel'][:50], r['ms'], r['frac_of_hbm_peak'])"
done
timeout 600 python tools/bench_hbm.py --only=agg --search > $OUT/hbm_search.json 2> $OUT/hbm_search.err; echo "search exit $?"
python -c "
import json
d=json.load(open('$OUT/hbm_search.json'))
print(d.get('neighbour_search_runs_ms'))
for r in d['rows'][1:]: print(r['kernel'][:60], r['ms'], r['frac_of_hbm_peak'])"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'nb_cell_fill|nb_cell_count' -c 2 -o $OUT/nb_full -f \
    python tools/bench_hbm.py --only=agg --search > $OUT/ncu_nb.log 2>&1; echo "ncu nb exit $?"
ls -la $OUT
