"""Two-rank NCCL test of pigs_b200.sharded.ShardedGaussianSampler on real GPUs (skipped on a one-GPU box): each rank
runs the CUDA product on its shard of the samples; outputs must equal the oracle's rows of the unsharded problem and
the all-reduced gradients the oracle's full-batch gradients (the oracle is the checker only)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from pigs_b200.sharded import ShardedGaussianSampler, shard_range
        from tests.fixtures import synthetic
        from tests.helpers import upstream
        rec = synthetic(3000, 20001, seed=23)
        N = rec["samples"].shape[0]
        lo, hi = shard_range(N, rank, world)
        d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
        d["samples"] = d["samples"][lo:hi].contiguous()
        for k in ("means", "values", "conics"):
            d[k].requires_grad_(True)
        s = ShardedGaussianSampler(False)
        s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
        u = s.sample_gaussians()
        du = s.sample_gaussians_derivative()
        ddu = s.sample_gaussians_laplacian()
        gs = [torch.from_numpy(upstream(rec, fn)[lo:hi]).cuda() for fn in range(3)]
        ((u * gs[0]).sum() + (du * gs[1]).sum() + (ddu * gs[2]).sum()).backward()
        torch.cuda.synchronize()
        np.savez(os.path.join(out_dir, "rank%d.npz" % rank), lo=lo, hi=hi, pairs=s.pair_count(),
                 u=u.detach().cpu().numpy(), du=du.detach().cpu().numpy(), ddu=ddu.detach().cpu().numpy(),
                 dmeans=d["means"].grad.cpu().numpy(), dvalues=d["values"].grad.cpu().numpy(),
                 dconics=d["conics"].grad.cpu().numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(600)
def test_two_gpu_sharding_matches_oracle(tmp_path):
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from oracle import pigs_oracle as po
    from tests.fixtures import synthetic
    from tests.helpers import TOL, rel_err, upstream
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    rec = synthetic(3000, 20001, seed=23)
    st = po.preprocess(rec["means"], rec["covariances"], rec["samples"])
    full = [po.forward(fn, st, rec["means"], rec["values"], rec["conics"], rec["samples"]) for fn in range(3)]
    want = [np.zeros_like(rec[k], dtype=np.float64) for k in ("means", "values", "conics")]
    for fn in range(3):
        for acc, part in zip(want, po.backward(fn, st, rec["means"], rec["values"], rec["conics"], rec["samples"],
                                               upstream(rec, fn))):
            acc += part.reshape(acc.shape)
    parts = [np.load(os.path.join(str(tmp_path), "rank%d.npz" % r)) for r in range(world)]
    assert sum(int(p["pairs"]) for p in parts) == po.pair_count(st)  # the global pair set, split by samples
    for p in parts:
        lo, hi = int(p["lo"]), int(p["hi"])
        for k, key in enumerate(("u", "du", "ddu")):
            assert rel_err(p[key], full[k][lo:hi]) <= TOL, key
        for key, w in zip(("dmeans", "dvalues", "dconics"), want):
            assert rel_err(p[key].reshape(w.shape), w) <= TOL, key
    for key in ("dmeans", "dvalues", "dconics"):  # every rank holds the same all-reduced gradient
        np.testing.assert_array_equal(parts[0][key], parts[1][key])


def _agg_worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from pigs_b200.sharded import ShardedGaussianSampler
        from tests.fixtures import AGG_RECIPES
        from tests.helpers import AGG_KEYS, agg_upstream
        rec = AGG_RECIPES["agg_model"]()
        d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
        s = ShardedGaussianSampler(False)
        s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
        s.preprocess_aggregate()
        t = [d[k].clone().requires_grad_(True) for k in AGG_KEYS]
        out = s.aggregate_neighbors(*t)
        out.backward(torch.from_numpy(agg_upstream(rec)).cuda())
        torch.cuda.synchronize()
        np.savez(os.path.join(out_dir, "agg_rank%d.npz" % rank), rows=np.array(s.row_range), slots=s.indices.numel(),
                 out=out.detach().cpu().numpy(), **{"grad_" + k: x.grad.cpu().numpy() for k, x in zip(AGG_KEYS, t)})
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(600)
def test_two_gpu_row_sharded_aggregation_matches_oracle(tmp_path):
    """C5 on two GPUs: each rank builds the neighbour lists of half of the query Gaussians; the all-reduced output and
    gradients equal the oracle's unsharded ones on both ranks."""
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from oracle import pigs_oracle as po
    from tests.fixtures import AGG_RECIPES
    from tests.helpers import AGG_KEYS, TOL, agg_upstream, oracle_state, rel_err
    world = 2
    mp.spawn(_agg_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    rec = AGG_RECIPES["agg_model"]()
    st = oracle_state(rec)
    idx, rng, dists, dens, inv = po.neighbors(rec["means"], rec["conics"], st["radii"])
    a = [rec[k] for k in AGG_KEYS]
    w, e, f, out = po.aggregate_forward(*a, idx, rng, dists, dens, inv)
    want = po.aggregate_backward(*a, idx, rng, dists, dens, w, e, f, inv, agg_upstream(rec))
    parts = [np.load(os.path.join(str(tmp_path), "agg_rank%d.npz" % r)) for r in range(world)]
    assert sum(int(p["slots"]) for p in parts) == idx.shape[0]   # the CSR, split by query Gaussian
    for p in parts:
        assert rel_err(p["out"], out) <= TOL
        for key, wv in zip(AGG_KEYS, want):
            assert rel_err(p["grad_" + key], wv) <= TOL, key
    for key in ["out"] + ["grad_" + k for k in AGG_KEYS]:
        np.testing.assert_array_equal(parts[0][key], parts[1][key])


def _rows_worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from pigs_b200.sharded import ShardedGaussianSampler
        from tests.fixtures import AGG_RECIPES
        from tests.helpers import AGG_KEYS, agg_upstream
        from tests.test_sharded_cpu import _local_csr, _window_csr
        rec = AGG_RECIPES["agg_model"]()
        P = rec["means"].shape[0]
        d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
        s = ShardedGaussianSampler(False)
        s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
        res = {}
        for tag in ("search", "local"):
            lo, hi = (0, P // 2 + 3) if rank == 0 else (P // 2 + 3, P)  # unequal windows
            if tag == "search":   # the reference's own neighbour rule: wide strips; forced onto the collective (dense) path
                s.preprocess_aggregate(row_range=(lo, hi), dense_fraction=0.2)
            else:                 # spatially local lists -> point-to-point halo exchange
                csr = _window_csr(*_local_csr(P, 16, seed=6), lo, hi)
                s.preprocess_aggregate(row_range=(lo, hi), csr=tuple(torch.from_numpy(x).cuda() for x in csr))
            t = {k: d[k].clone() for k in AGG_KEYS}
            for k in ("features", "queries", "keys"):
                t[k] = t[k][lo:hi].contiguous()      # a rank holds only its own rows of the row-indexed inputs
            for k in AGG_KEYS:
                t[k].requires_grad_(True)
            out = s.aggregate_neighbors_rows(*[t[k] for k in AGG_KEYS])
            out.backward(torch.from_numpy(agg_upstream(rec)[lo:hi]).cuda())
            torch.cuda.synchronize()
            # the multi-head node with ONE head must reproduce the per-head call (same kernels, one exchange each way)
            t2 = {k: t[k].detach().clone().requires_grad_(True) for k in AGG_KEYS}
            (o2,) = s.aggregate_neighbors_rows_heads(t2["features"], t2["frequencies"],
                                                     [(t2["transform"], t2["queries"], t2["keys"], t2["distance_transform"])])
            o2.backward(torch.from_numpy(agg_upstream(rec)[lo:hi]).cuda())
            torch.cuda.synchronize()
            assert torch.equal(o2.detach(), out.detach())
            for k in AGG_KEYS:
                den = float(t[k].grad.abs().max()) + 1e-30
                assert float((t2[k].grad - t[k].grad).abs().max()) <= 1e-5 * den, (tag, k)
            res.update({tag + "_out": out.detach().cpu().numpy(), tag + "_dense": int(s.row_plan.dense),
                        **{tag + "_grad_" + k: t[k].grad.cpu().numpy() for k in AGG_KEYS}})
        np.savez(os.path.join(out_dir, "rows_rank%d.npz" % rank), **res)
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(600)
def test_two_gpu_row_parallel_aggregation_matches_oracle(tmp_path):
    """Row-parallel C5 on two GPUs (ShardedGaussianSampler.aggregate_neighbors_rows): each rank holds its own rows of
    features / queries / keys, the halo rows travel point to point (or, for non-local lists, by all-gather), and each
    rank gets its own rows of the output and of the gradients; concatenated they equal the oracle's unsharded run."""
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from oracle import pigs_oracle as po
    from tests.fixtures import AGG_RECIPES
    from tests.helpers import AGG_KEYS, TOL, agg_upstream, oracle_state, rel_err
    from tests.test_sharded_cpu import _local_csr, _window_csr
    world = 2
    mp.spawn(_rows_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    rec = AGG_RECIPES["agg_model"]()
    P = rec["means"].shape[0]
    st = oracle_state(rec)
    a = [rec[k] for k in AGG_KEYS]
    parts = [np.load(os.path.join(str(tmp_path), "rows_rank%d.npz" % r)) for r in range(world)]
    for tag in ("search", "local"):
        csr = po.neighbors(rec["means"], rec["conics"], st["radii"]) if tag == "search" else \
            _window_csr(*_local_csr(P, 16, seed=6), 0, P)
        w, e, f, out = po.aggregate_forward(*a, *csr)
        want = po.aggregate_backward(*a, csr[0], csr[1], csr[2], csr[3], w, e, f, csr[4], agg_upstream(rec))
        assert int(parts[0][tag + "_dense"]) == int(parts[1][tag + "_dense"])  # one decision, taken from the gathered table
        if tag == "local":
            assert int(parts[0][tag + "_dense"]) == 0                           # halo rows travel point to point
        assert rel_err(np.concatenate([p[tag + "_out"] for p in parts]), out) <= TOL
        for key, wv in zip(AGG_KEYS, want):
            if key in ("features", "queries", "keys"):
                got = np.concatenate([p[tag + "_grad_" + key] for p in parts])
                assert rel_err(got, wv) <= TOL, (tag, key)
            else:
                for p in parts:
                    assert rel_err(p[tag + "_grad_" + key], wv) <= TOL, (tag, key)


@pytest.mark.gpu
def test_pair_weights_match_oracle_and_balance_bands():
    """GaussianSampler.pair_weights(): per-sample pair counts equal the oracle's (Gaussians binned to the sample's tile),
    and windows cut from them carry equal work where an even split of grid-ordered means does not."""
    import numpy as np
    import torch

    from oracle import pigs_oracle as po
    from pigs_b200 import GaussianSampler
    from pigs_b200.sharded import balanced_ranges, shard_range
    from tests.fixtures import synthetic
    rec = synthetic(20000, 16, seed=3)
    rec["samples"] = rec["means"].copy()                      # model_pn.py:648: the means are the samples
    st = po.preprocess(rec["means"], rec["covariances"], rec["samples"])
    d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
    s = GaussianSampler(False)
    s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
    w = s.pair_weights()
    assert int(w.sum()) == s.pair_count()
    blocks = st["blocks"]
    ng = (st["ranges"][:blocks, 1].astype(np.int64) - st["ranges"][:blocks, 0].astype(np.int64))
    want = np.zeros(rec["samples"].shape[0], np.int64)
    for t in range(blocks):
        a, b = int(st["sample_ranges"][t, 0]), int(st["sample_ranges"][t, 1])
        want[st["sample_list"][a:b].astype(np.int64)] = ng[t]
    np.testing.assert_array_equal(w.cpu().numpy(), want)
    world = 8
    bal = [int(w[lo:hi].sum()) for lo, hi in balanced_ranges(w, world)]
    even = [int(w[slice(*shard_range(w.numel(), r, world))].sum()) for r in range(world)]
    assert max(bal) <= max(even) and max(bal) - min(bal) <= 2 * int(w.max())
