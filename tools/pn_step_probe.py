"""Dev tool: one model_pn.py-style step at the reference's OWN scale (main_pn.py: nx = ny = 20 -> 400 Gaussians; also
900 / 2 500 / 10 000), through the public API: preprocess(samples = means) -> u, grad u, Hessian -> preprocess_aggregate
(the real neighbour search) -> 2 heads of aggregate_neighbors, then the backward of a scalar loss through both.
Ours next to the UNMODIFIED reference extension (oracle/_ref/_C.so) where it is on the box.  One JSON line per size."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from diff_gaussian_sampling import GaussianSampler  # noqa: E402
from pigs_b200 import _lib  # noqa: E402
from tests import refcuda  # noqa: E402
from tests.fixtures import agg_model  # noqa: E402
from tests.helpers import AGG_KEYS  # noqa: E402

L = _lib.lib()
HEADS = 2


def ours_step(s, d, heads):
    s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["means"])
    u, ux, uxx = s.sample_gaussians(), s.sample_gaussians_derivative(), s.sample_gaussians_laplacian()
    s.preprocess_aggregate()
    loss = u.sum() + ux.sum() + uxx.sum()
    for h in heads:
        loss = loss + s.aggregate_neighbors(*h).sum()
    loss.backward()
    return loss


def ref_step(r, C, d, heads):
    r.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["means"])
    outs = [r.forward(k) for k in range(3)]
    csr = r.preprocess_aggregate()
    res = []
    for h in heads:
        w, e, f, out = C.aggregate_neighbors(*h, *csr, False)
        res.append(C.aggregate_neighbors_backward(*h, csr[0], csr[1], csr[2], csr[3], w, e, f, csr[4], torch.ones_like(out), False))
    grads = [r.backward(k, torch.ones_like(outs[k])) for k in range(3)]
    return outs, res, grads


def timed(fn, steps, warmup):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    n0 = L.pigs_launch_count()
    ev[0].record()
    for i in range(steps):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    per = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(steps))
    return per[len(per) // 2], (L.pigs_launch_count() - n0) / steps


for P in [int(x) for x in os.environ.get("PN_SIZES", "400,900,2500,4096,10000").split(",")]:
    rec = agg_model(P, seed=21)
    d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
    for k in ("means", "values", "conics"):
        d[k].requires_grad_(True)
    heads = []
    for h in range(HEADS):
        heads.append([d[k].clone().requires_grad_(True) for k in AGG_KEYS])
    s = GaussianSampler(False)
    ms, launches = timed(lambda: ours_step(s, d, heads), 30, 10)
    row = {"P": P, "ours_ms": round(ms, 4), "ours_launches": launches, "neighbour_slots": int(s.indices.numel())}
    if refcuda.available() and P <= 10000 and not os.environ.get("PN_NO_REF"):
        r = refcuda.RefSampler(False)
        dd = {k: v.detach() for k, v in d.items()}
        hh = [[t.detach() for t in h] for h in heads]
        rms, _ = timed(lambda: ref_step(r, r.C, dd, hh), 5 if P > 2500 else 10, 2)
        row["ref_cuda_ms"] = round(rms, 4)
        row["speedup"] = round(rms / ms, 2)
    print(json.dumps(row), flush=True)
