"""Dev tool: device-resident step (preprocess + fused u / grad u / Hessian forward + backward) of small problems, with the
small-problem mode (one evaluation kernel, pigs_set_small_problem_pairs) on and off.  Prints one JSON line."""
import json
import sys
import os

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pigs_b200 import _C, _lib  # noqa: E402
from tests.fixtures import synthetic  # noqa: E402
from tests.helpers import upstream  # noqa: E402

L = _lib.lib()
SIZES = [(1000, 10000, 1), (100, 4096, 2), (100, 16384, 2), (100, 40000, 2), (400, 4096, 2), (900, 10000, 2), (2000, 12000, 2), (3000, 20000, 2),
         (4000, 40000, 2), (5000, 50000, 2)]


def step(d, ups):
    nr, bb, sbb, rg, srg, radii = _C.preprocess_gaussians(d["means"], d["values"], d["covariances"], d["conics"], d["samples"], False)
    a = (d["means"], d["values"], d["conics"], d["samples"], nr)
    outs = _C.sample_forward(7, *a, bb, sbb, rg, srg, False)
    return nr, rg.numel() // 8 - 1, _C.sample_backward(7, *a, [ups[0], ups[1], ups[2], None], bb, sbb, rg, srg, False)


def timed(d, ups, steps=40, warmup=10):
    for _ in range(warmup):
        step(d, ups)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n0 = L.pigs_launch_count()
    a.record()
    for _ in range(steps):
        nr, blocks, _ = step(d, ups)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / steps, (L.pigs_launch_count() - n0) / steps, nr, blocks


rows = []
for P, N, D in SIZES:
    rec = synthetic(P, N, seed=3, D=D)
    d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
    ups = [torch.from_numpy(upstream(rec, fn)).cuda() for fn in range(3)]
    row = {"P": P, "N": N, "D": D}
    for name, pairs in (("three_class", 0), ("single_class", 1 << 40)):
        L.pigs_set_small_problem_pairs(pairs)
        ms, launches, nr, blocks = timed(d, ups)
        row[name] = {"ms_per_step": round(ms, 4), "launches": launches}
    row["num_rendered"], row["blocks"] = nr, blocks
    row["estimated_pairs"] = nr * N // blocks
    rows.append(row)
    print(json.dumps(row), flush=True)
