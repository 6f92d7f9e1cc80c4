#!/bin/bash
set -u
TAG=${1:-rXX}
OUT=gpurun_out/$TAG
mkdir -p $OUT
for w in 16 24 32 64 128 422; do
  PIGS_A16_FWD_CTAS_PER_SM=$w timeout 300 python tools/bench_hbm.py --only=agg > $OUT/hbm_ctas$w.json 2> $OUT/hbm_ctas$w.err
  python -c "
import json
for r in json.load(open('$OUT/hbm_ctas$w.json'))['rows'][1:2]: print('ctas/SM $w', r['kernel'][:50], r['ms'], r['frac_of_hbm_peak'])"
done
