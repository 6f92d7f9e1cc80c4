#!/bin/bash
# Short GPU-box pass: parity tests, smoke, the bench line with its legs, then (optionally) the sanitizer pass.
# usage: tools/gpu_pass.sh <tag> [sanitize-seconds-per-invocation, 0 = skip]
set -u
TAG=${1:-rXX}
SAN=${2:-0}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $OUT/smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_gpu.log
tail -4 $OUT/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke exit $?" >> $OUT/smoke.log
tail -2 $OUT/smoke.log
timeout 900 python bench.py > $OUT/bench_c3.json 2> $OUT/bench_c3.err; echo "bench exit $?"
tail -c 6000 $OUT/bench_c3.json
tail -5 $OUT/bench_c3.err
timeout 600 python bench.py --impl reference > $OUT/bench_reference_arm.json 2> $OUT/bench_ref.err; echo "ref arm exit $?"
if [ "$SAN" != "0" ]; then tools/sanitize.sh $TAG $SAN; fi
ls -la $OUT
