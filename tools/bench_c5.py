"""Config C5 (BASELINE.json configs[4]: model_pn.py neighbour aggregation + PDE step, 1M Gaussians, 16 neighbours) as the
sequence of library calls one `Model.forward` + loss backward makes (model_pn.py:644-700), timed stage by stage:
  preprocess(samples = means) -> fused u / grad u / Hessian at the means (no grad, model_pn.py:647-651)
  -> aggregation forward + backward on a 16-regular CSR (the reference's own neighbour rule gives ~1.6 K
     neighbours per Gaussian at this P; `_C.aggregate_neighbors` takes the CSR as plain inputs)
The MLP between them is torch (out of scope).  python tools/bench_c5.py > profiles/rNN_c5_stages.json
Under torchrun (WORLD_SIZE > 1) the same step runs row-sharded: every rank samples at ITS contiguous range
of the means, owns the neighbour rows of that range, and the step carries three all-reduces (tile-grid bounds, the
[P, L] aggregation output, the packed aggregation gradients); rank 0 prints one JSON object with the max-over-ranks
device time.  torchrun --nproc-per-node N tools/bench_c5.py > profiles/rNN_c5_nN.json"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
from bench_hbm import regular_csr, timed  # noqa: E402
from pigs_b200 import GaussianSampler, _C  # noqa: E402
from tests.fixtures import agg_params, synthetic  # noqa: E402


def main_sharded(world):
    import torch.distributed as dist

    import pigs_b200
    from pigs_b200.sharded import AggregateReduce, ShardedGaussianSampler, balanced_ranges, shard_range
    rank, local = int(os.environ["RANK"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    P, k = 1_000_000, 16
    rec = synthetic(P, 16, seed=0)
    d = {kk: torch.from_numpy(v).cuda() for kk, v in rec.items()}
    s = ShardedGaussianSampler(False)
    if os.environ.get("PIGS_C5_EVEN") == "1":  # (a --even flag would be eaten by torchrun's own prefix matching)
        lo, hi = shard_range(P, rank, world)
        split = "even row counts"
    else:
        # one binning pass over ALL means (set-up, not timed: the windows only change when the Gaussians move between
        # tiles) -> per-row pair counts -> windows of equal evaluation work
        s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["means"])
        lo, hi = balanced_ranges(s.pair_weights(), world)[rank]
        split = "equal pair counts (sharded.balanced_ranges over GaussianSampler.pair_weights)"
    mine = d["means"][lo:hi].contiguous()
    idx, _ = regular_csr(rec["means"], k)
    indices = torch.from_numpy(idx[lo * k:hi * k]).cuda()
    ranges = ((torch.arange(1, P + 1, device="cuda") - lo).clamp(0, hi - lo) * k).contiguous()
    i_of = torch.arange(lo, hi, device="cuda").repeat_interleave(k)
    s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], mine)
    X = d["means"][indices] - d["means"][i_of]
    X = torch.where(X.abs() > 1, X - 2 * torch.sign(X), X)
    dists = (X / (0.333 * s.radii[i_of] + 1e-6).unsqueeze(1)).contiguous()
    cj = d["conics"][indices]
    dens = torch.exp(-0.5 * (cj[:, 0] * X[:, 0] ** 2 + cj[:, 2] * X[:, 1] ** 2) - cj[:, 1] * X[:, 0] * X[:, 1]).contiguous()
    inv_total = torch.full((P,), 1e6, device="cuda")
    inv_total[lo:hi] = 1.0 / (dens.view(hi - lo, k).sum(1) + 1e-6)
    ap = {kk: torch.from_numpy(v).cuda().requires_grad_(True) for kk, v in agg_params(P, 16, 16, 25, 2, seed=100).items()}
    a = [ap[n] for n in ("features", "transform", "queries", "keys", "frequencies", "distance_transform")]
    g = torch.randn(P, 16, device="cuda")
    red = AggregateReduce()
    mine_pairs = torch.zeros(world, device="cuda", dtype=torch.int64)
    mine_pairs[rank] = s.pair_count()
    dist.all_reduce(mine_pairs)
    per_rank, pairs = mine_pairs.tolist(), mine_pairs.sum()

    def step():
        s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], mine)
        with torch.no_grad():
            s.sample_fused(1 | 2 | 16)
        o = pigs_b200.aggregate_neighbors(*a, indices, ranges, dists, dens, inv_total, False, reduce=red)
        for t in a:
            t.grad = None
        o.backward(g)

    def sampling():
        s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], mine)
        with torch.no_grad():
            s.sample_fused(1 | 2 | 16)

    res = {}
    for name, fn in (("step (preprocess + u/grad u/Laplacian at own means + aggregate fwd + bwd, 3 all-reduces)", step),
                     ("sampling part alone", sampling)):
        for _ in range(3):
            fn()
        ts = []
        for _ in range(5):
            torch.cuda.synchronize()
            dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1)], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ts.append(float(t.item()))
        res[name] = float(np.median(ts))
    if rank == 0:
        print(json.dumps({"gpu": torch.cuda.get_device_name(0), "n_gpus": world, "gaussians": P, "neighbours": k,
                          "row_split": split, "rows_rank0": hi - lo, "pairs_per_rank": per_rank,
                          "pairs_of_the_sampling_pass_all_ranks": int(pairs.item()),
                          "timing": "CUDA events on the compute stream, barrier + synchronize either side, max over ranks, median of 5",
                          "ms": res}, indent=1))
    dist.destroy_process_group()


def main():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        return main_sharded(world)
    torch.cuda.set_device(0)
    P = 1_000_000
    rec = synthetic(P, 16, seed=0)
    d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
    s = GaussianSampler(False)
    pre = lambda: s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["means"])
    t_pre = timed(pre)
    pairs = s.pair_count()
    with torch.no_grad():
        t_fwd = timed(lambda: s.sample_fused(7), it=5, warm=2)
        t_pde = timed(lambda: s.sample_fused(1 | 2 | 16), it=5, warm=2)
    k = 16
    idx, rng = regular_csr(rec["means"], k)
    indices, ranges = torch.from_numpy(idx).cuda(), torch.from_numpy(rng).cuda()
    i_of = torch.arange(P, device="cuda").repeat_interleave(k)
    X = d["means"][indices] - d["means"][i_of]
    X = torch.where(X.abs() > 1, X - 2 * torch.sign(X), X)
    dists = (X / (0.333 * s.radii[i_of] + 1e-6).unsqueeze(1)).contiguous()
    cj = d["conics"][indices]
    dens = torch.exp(-0.5 * (cj[:, 0] * X[:, 0] ** 2 + cj[:, 2] * X[:, 1] ** 2) - cj[:, 1] * X[:, 0] * X[:, 1]).contiguous()
    inv_total = (1.0 / (dens.view(P, k).sum(1) + 1e-6)).contiguous()
    ap = {kk: torch.from_numpy(v).cuda() for kk, v in agg_params(P, 16, 16, 25, 2, seed=100).items()}
    a = (ap["features"], ap["transform"], ap["queries"], ap["keys"], ap["frequencies"], ap["distance_transform"])
    t_af = timed(lambda: _C.aggregate_neighbors(*a, indices, ranges, dists, dens, inv_total, False))
    w, e, f, o = _C.aggregate_neighbors(*a, indices, ranges, dists, dens, inv_total, False)
    g = torch.randn_like(o)
    t_ab = timed(lambda: _C.aggregate_neighbors_backward(*a, indices, ranges, dists, dens, w, e, f, inv_total, g, False))
    out = {"gpu": torch.cuda.get_device_name(0), "gaussians": P, "samples": "the means (model_pn.py:648)",
           "pairs_of_the_sampling_pass": pairs,
           "ms": {"preprocess": t_pre, "sample u + grad u + Hessian (no grad)": t_fwd,
                  "sample u + grad u + Laplacian (no grad, PIGS_FN_LAPL)": t_pde,
                  "aggregate forward (16-regular CSR, one head)": t_af, "aggregate backward": t_ab},
           "step_ms_hessian": t_pre + t_fwd + t_af + t_ab, "step_ms_laplacian": t_pre + t_pde + t_af + t_ab}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
