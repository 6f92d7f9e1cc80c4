#!/bin/bash
# Multi-GPU pass on one box: NCCL parity tests (2 ranks) and the bench line with its c5 leg at N ranks.
# usage: tools/gpu_multi.sh <tag> <N> [test: 1|0]
set -u
TAG=${1:-rXX}
N=${2:-2}
T=${3:-1}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=index,name --format=csv > $OUT/smi.txt 2>&1
if [ "$T" = "1" ]; then
  timeout 900 python -m pytest tests/test_gpu_sharded.py tests/test_compiled_binding.py -m gpu -x -q > $OUT/pytest_multi.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_multi.log
  tail -5 $OUT/pytest_multi.log
fi
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus $N --steps 5 --warmup 3 > $OUT/bench_c3_n$N.json 2> $OUT/bench_n$N.err; echo "bench N=$N exit $?"
tail -5 $OUT/bench_n$N.err
python - <<PY
import json
d=json.load(open("$OUT/bench_c3_n$N.json"))
print("C3 N=$N", d["ms_per_step"], d["kernel_ms"]["per_rank_fwd_bwd_ms"], "e2e", d["e2e"]["ms_per_step"])
x=d.get("c5",{}); print("c5", {k:x.get(k) for k in ("ms_per_step","sampling_ms","aggregation_ms","gpu_launches_per_step","error","exchange","halo_rows_rank0","rows_per_rank")})
PY
ls -la $OUT
