#!/bin/bash
# One GPU-box pass: parity tests, bench (no profiler), then ncu launch list + one --set full capture.
# usage: tools/gpu_round.sh <tag>
set -u
TAG=${1:-rXX}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $OUT/smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_gpu.log
tail -3 $OUT/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke exit $?" >> $OUT/smoke.log
tail -2 $OUT/smoke.log
timeout 900 python bench.py > $OUT/bench_c3.json 2> $OUT/bench_c3.err; echo "bench exit $?"
tail -c 3000 $OUT/bench_c3.json
timeout 600 python bench.py --impl reference > $OUT/bench_reference_arm.json 2> $OUT/bench_ref.err; echo "ref arm exit $?"
timeout 600 python tools/quick_perf.py --big > $OUT/quick_perf.log 2>&1; echo "quick exit $?"
cat $OUT/quick_perf.log | tail -12
timeout 600 python tools/bench_hbm.py > $OUT/hbm_kernels.json 2> $OUT/hbm.err; echo "hbm exit $?"
timeout 300 python tools/bench_dense.py > $OUT/dense.json 2> $OUT/dense.err; echo "dense exit $?"
timeout 300 python tools/bench_c5.py > $OUT/c5_stages.json 2> $OUT/c5.err; echo "c5 exit $?"
# launch list of the same bench command (per-launch times are cold-cache and serialised)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/launches_c3.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > $OUT/ncu_launch.log 2>&1; echo "ncu launch exit $?"
# full capture of the six evaluation kernels of one C3 step (the bench workload itself)
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:eval_ -c 6 -o $OUT/eval_c3_full -f \
    python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e > $OUT/ncu_full.log 2>&1; echo "ncu full exit $?"
ls -la $OUT
