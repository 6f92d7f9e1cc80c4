"""Dev tool: forward + backward of a small problem (C1: 1-D 1K x 10K; 2-D 900 x 10K) over a fixed binning, eager against
replayed from a CUDA graph (tests/test_graph_capture.py holds the replay to the oracle)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pigs_b200 import _C  # noqa: E402
from tests.fixtures import synthetic  # noqa: E402
from tests.helpers import upstream  # noqa: E402


def timed(fn, steps=200, warmup=20):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / steps


for P, N, D in ((1000, 10000, 1), (900, 10000, 2), (3000, 20000, 2)):
    rec = synthetic(P, N, seed=3, D=D)
    d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
    ups = [torch.from_numpy(upstream(rec, fn)).cuda() for fn in range(3)]
    nr, bb, sbb, rg, srg, _ = _C.preprocess_gaussians(d["means"], d["values"], d["covariances"], d["conics"], d["samples"], False)
    a = (d["means"], d["values"], d["conics"], d["samples"], nr)

    def step():
        _C.sample_forward(7, *a, bb, sbb, rg, srg, False)
        _C.sample_backward(7, *a, [ups[0], ups[1], ups[2], None], bb, sbb, rg, srg, False)

    eager = timed(step)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3):
            step()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side):
        step()
    replay = timed(g.replay)
    print(json.dumps({"P": P, "N": N, "D": D, "fwd_bwd_eager_ms": round(eager, 4), "fwd_bwd_graph_replay_ms": round(replay, 4)}), flush=True)
