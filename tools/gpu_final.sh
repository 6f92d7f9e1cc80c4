#!/bin/bash
# Final pass of a round on one GPU: parity tests, smoke, bench + reference arm, variants / HBM / dense tables, then the
# profiler passes (launch list + `ncu --set full` of the evaluation kernels of the bench workload).
# usage: tools/gpu_final.sh <tag>
set -u
TAG=${1:-rXX}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $OUT/smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_gpu.log
tail -4 $OUT/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke exit $?" >> $OUT/smoke.log
tail -2 $OUT/smoke.log
timeout 900 python bench.py > $OUT/bench_c3.json 2> $OUT/bench_c3.err; echo "bench exit $?"
timeout 600 python bench.py --impl reference > $OUT/bench_reference_arm.json 2> $OUT/bench_ref.err; echo "ref arm exit $?"
timeout 600 python tools/bench_variants.py > $OUT/variants.json 2> $OUT/variants.err; echo "variants exit $?"
timeout 600 python tools/bench_hbm.py > $OUT/hbm_kernels.json 2> $OUT/hbm.err; echo "hbm exit $?"
timeout 300 python tools/bench_dense.py > $OUT/dense.json 2> $OUT/dense.err; echo "dense exit $?"
timeout 300 python tools/fwd_shard_probe.py > $OUT/fwd_shard_probe.json 2> $OUT/probe.err; echo "probe exit $?"
for w in C3 C1 shard8; do timeout 120 python tools/chain_timing.py $w > $OUT/chain_timing_$w.json 2>> $OUT/chain.err; done
python - <<PY
import json
d=json.load(open("$OUT/bench_c3.json"))
print("C3", d["ms_per_step"], d["kernel_ms"]["eval_fwd"], d["kernel_ms"]["eval_bwd"], d["kernel_ms"]["other_per_step"], "e2e", d["e2e"]["ms_per_step"], "launches", d["gpu_launches"], "traffic", d["roofline"]["traffic"])
for k in ("c5","c1","c2","c4"):
    x=d.get(k,{}); print(k, x.get("ms_per_step"), x.get("sampling_ms"), x.get("aggregation_ms"), x.get("gpu_launches_per_step"), (x.get("e2e") or {}).get("ms_per_step"), x.get("error"))
print("ref_cuda", d.get("ref_cuda",{}).get("speedup"), d.get("ref_cuda",{}).get("error"))
for r in json.load(open("$OUT/hbm_kernels.json"))["rows"]: print(r["kernel"][:60], r["ms"], r["frac_of_hbm_peak"])
PY
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_c3.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-legs > $OUT/ncu_launch.log 2>&1; echo "ncu launch exit $?"
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:eval_ -c 6 -o $OUT/eval_c3_full -f \
    python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --no-legs > $OUT/ncu_full.log 2>&1; echo "ncu eval exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:agg_ --launch-skip 8 -c 4 -o $OUT/agg_full -f \
    python tools/bench_hbm.py --only=agg > $OUT/ncu_agg.log 2>&1; echo "ncu agg exit $?"
ls -la $OUT
