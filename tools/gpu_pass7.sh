#!/bin/bash
# tests + dense path + HBM-bound kernels + bench line.  usage: tools/gpu_pass7.sh <tag>
set -u
TAG=${1:-rXX}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_gpu.log
tail -4 $OUT/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke exit $?" >> $OUT/smoke.log
tail -2 $OUT/smoke.log
timeout 300 python tools/bench_dense.py > $OUT/dense.json 2> $OUT/dense.err; echo "dense exit $?"
python -c "
import json
for r in json.load(open('$OUT/dense.json'))['rows']: print(r['workload'], round(r['fwd_ms'],3), round(r['fwd_bwd_ms'],3))" 2>/dev/null || tail -3 $OUT/dense.err
timeout 600 python tools/bench_hbm.py > $OUT/hbm_kernels.json 2> $OUT/hbm.err; echo "hbm exit $?"
python -c "
import json
for r in json.load(open('$OUT/hbm_kernels.json'))['rows']: print(r['kernel'][:70], r['ms'], r['frac_of_hbm_peak'])"
timeout 600 python tools/bench_variants.py > $OUT/variants.json 2> $OUT/variants.err; echo "variants exit $?"
timeout 900 python bench.py > $OUT/bench_c3.json 2> $OUT/bench_c3.err; echo "bench exit $?"
python - <<PY
import json
d=json.load(open("$OUT/bench_c3.json"))
print("C3", d["ms_per_step"], d["kernel_ms"]["eval_fwd"], d["kernel_ms"]["eval_bwd"], d["kernel_ms"]["other_per_step"], "e2e", d["e2e"]["ms_per_step"], "launches", d["gpu_launches"])
for k in ("c5","c1","c2","c4"):
    x=d.get(k,{}); print(k, x.get("ms_per_step"), x.get("sampling_ms"), x.get("aggregation_ms"), x.get("gpu_launches_per_step"), (x.get("e2e") or {}).get("ms_per_step"), x.get("error"))
PY
ls -la $OUT
