#!/bin/bash
# Profiling pass (one GPU): launch list of the bench command, then `ncu --set full` captures of the evaluation kernels
# (C3, the bench workload itself), the aggregation kernels (P = 1M, 16-regular CSR) and the binning / staging kernels.
# Numbers printed by runs under ncu are never bench values.   usage: tools/gpu_profile.sh <tag>
set -u
TAG=${1:-rXX}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_aggregate.py tests/test_gpu_parity.py tests/test_densify.py -m gpu -x -q > $OUT/pytest_subset.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_subset.log
tail -3 $OUT/pytest_subset.log
timeout 600 python tools/bench_hbm.py > $OUT/hbm_kernels.json 2> $OUT/hbm.err; echo "hbm exit $?"
python -c "
import json
for r in json.load(open('$OUT/hbm_kernels.json'))['rows']: print(r['kernel'][:70], r['ms'], r['frac_of_hbm_peak'])"
# launch list of the same bench command (per-launch times are cold-cache and serialised)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_c3.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-legs > $OUT/ncu_launch.log 2>&1; echo "ncu launch exit $?"
# full capture of the six evaluation kernels of one C3 step (the bench workload itself)
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:eval_ -c 6 -o $OUT/eval_c3_full -f \
    python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --no-legs > $OUT/ncu_full.log 2>&1; echo "ncu eval exit $?"
# aggregation kernels (the C5 leg's kernels at P = 1M)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:agg_ -c 4 -o $OUT/agg_full -f \
    python tools/bench_hbm.py > $OUT/ncu_agg.log 2>&1; echo "ncu agg exit $?"
# binning / staging kernels of one C3 preprocess + forward + backward
timeout 900 ncu --set full --clock-control none -k regex:'radix_|sample_key|emit_|flag_|pack_|count_|scan_|bounds_|worklist' -c 24 -o $OUT/binning_full -f \
    python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --no-legs > $OUT/ncu_bin.log 2>&1; echo "ncu binning exit $?"
ls -la $OUT
