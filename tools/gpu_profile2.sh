#!/bin/bash
# binning / staging kernels only: ncu --set full of one C3 preprocess + forward + backward.  usage: tools/gpu_profile2.sh <tag>
set -u
TAG=${1:-rXX}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 ncu --set full --clock-control none -k regex:'radix_|sample_key|emit_|flag_|pack_|count_|scan_|bounds_|worklist' -c 18 -o $OUT/binning_full -f \
    python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --no-legs > $OUT/ncu_bin.log 2>&1; echo "ncu binning exit $?"
ls -la $OUT
