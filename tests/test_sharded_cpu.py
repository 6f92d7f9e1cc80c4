"""world_size-2 gloo test (CPU) of the multi-GPU host logic in pigs_b200/sharded.py: contiguous sample shards,
the bounds all-reduce that keeps the tile grid global, and the packed-gradient all-reduce.  The per-rank compute
is done HERE by the CPU oracle (test infrastructure) — the product has no CPU path; what is checked is that the
sharding + collectives reproduce the single-process result: same tile grid and pair set, outputs equal to the
unsharded rows, gradients equal to the unsharded gradients after the all-reduce."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import pigs_oracle as po
        from pigs_b200.sharded import allreduce_bounds, allreduce_grads, shard_range
        from tests.fixtures import synthetic
        from tests.helpers import upstream
        rec = synthetic(1200, 6001, seed=21)  # odd N: uneven shards
        N, D = rec["samples"].shape
        lo, hi = shard_range(N, rank, world)
        mine = rec["samples"][lo:hi]
        mm = torch.from_numpy(np.concatenate([mine.min(0), mine.max(0)]).astype(np.float32))
        allreduce_bounds(mm, D)
        st = po.preprocess(rec["means"], rec["covariances"], mine, bounds=(mm[:D].numpy(), mm[D:].numpy()))
        flat = []
        outs = []
        for fn in range(3):
            outs.append(po.forward(fn, st, rec["means"], rec["values"], rec["conics"], mine))
            g = upstream(rec, fn)[lo:hi]
            flat.append(np.concatenate([x.reshape(-1) for x in po.backward(fn, st, rec["means"], rec["values"],
                                                                           rec["conics"], mine, g)]))
        packed = torch.from_numpy(np.sum(flat, 0).astype(np.float32))
        allreduce_grads(packed)
        np.savez(os.path.join(out_dir, "rank%d.npz" % rank), lo=lo, hi=hi, mm=mm.numpy(), grid=st["grid"],
                 pairs=po.pair_count(st), packed=packed.numpy(), u=outs[0], du=outs[1], ddu=outs[2])
    finally:
        dist.destroy_process_group()


def test_shard_range_covers_everything():
    from pigs_b200.sharded import shard_range
    for N in (0, 1, 7, 8, 4_000_000, 4_000_003):
        for world in (1, 2, 3, 8):
            spans = [shard_range(N, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == N
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


@pytest.mark.timeout(300)
def test_two_rank_sharding_matches_single_process(tmp_path):
    from oracle import pigs_oracle as po
    from tests.fixtures import synthetic
    from tests.helpers import TOL, rel_err, upstream
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    rec = synthetic(1200, 6001, seed=21)
    st = po.preprocess(rec["means"], rec["covariances"], rec["samples"])
    full = [po.forward(fn, st, rec["means"], rec["values"], rec["conics"], rec["samples"]) for fn in range(3)]
    want = np.sum([np.concatenate([x.reshape(-1).astype(np.float64) for x in
                                   po.backward(fn, st, rec["means"], rec["values"], rec["conics"], rec["samples"],
                                               upstream(rec, fn))]) for fn in range(3)], 0)
    parts = [np.load(os.path.join(str(tmp_path), "rank%d.npz" % r)) for r in range(world)]
    assert sum(int(p["pairs"]) for p in parts) == po.pair_count(st)  # same global pair set
    for p in parts:
        np.testing.assert_array_equal(p["grid"], st["grid"])
        np.testing.assert_array_equal(p["mm"][:2], st["min_b"])
        lo, hi = int(p["lo"]), int(p["hi"])
        for k, key in enumerate(("u", "du", "ddu")):
            np.testing.assert_array_equal(p[key], full[k][lo:hi])  # per-sample results do not depend on the shard
        assert rel_err(p["packed"], want) <= TOL
    np.testing.assert_array_equal(parts[0]["packed"], parts[1]["packed"])


def _agg_worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import pigs_oracle as po
        from pigs_b200.sharded import AggregateReduce, shard_range
        from tests.fixtures import AGG_RECIPES
        from tests.helpers import AGG_KEYS, agg_upstream, oracle_state
        rec = AGG_RECIPES["agg_model"]()
        st = oracle_state(rec)
        idx, rng, dists, dens, inv = po.neighbors(rec["means"], rec["conics"], st["radii"])
        P = rng.shape[0]
        lo, hi = shard_range(P, rank, world)
        # this rank's CSR: the rows [lo, hi) of the full one, every other row empty with the no-neighbour total
        cnt = np.diff(rng, prepend=0)
        starts = rng - cnt
        a0, b0 = int(starts[lo]), int(rng[hi - 1])
        cnt_r = np.zeros_like(cnt)
        cnt_r[lo:hi] = cnt[lo:hi]
        inv_r = np.full_like(inv, np.float32(1.0 / 1e-6))
        inv_r[lo:hi] = inv[lo:hi]
        csr = (idx[a0:b0].copy(), np.cumsum(cnt_r).astype(np.int64), dists[a0:b0].copy(), dens[a0:b0].copy(), inv_r)
        a = [rec[k] for k in AGG_KEYS]
        w, e, f, out = po.aggregate_forward(*a, *csr)
        red = AggregateReduce(windows=[shard_range(P, r, world) for r in range(world)])  # row-owned tensors all-gathered
        out_t = red.output(torch.from_numpy(np.ascontiguousarray(out, np.float32)))
        grads = po.aggregate_backward(*a, csr[0], csr[1], csr[2], csr[3], w, e, f, csr[4], agg_upstream(rec))
        gl = [torch.from_numpy(np.ascontiguousarray(g, np.float32)) for g in grads]
        got = red.gradients(gl)
        np.savez(os.path.join(out_dir, "agg%d.npz" % rank), out=out_t.numpy(),
                 **{"grad_" + k: g.numpy() for k, g in zip(AGG_KEYS, got)})
    finally:
        dist.destroy_process_group()


def test_pack_unpack_grads_roundtrip():
    from pigs_b200.sharded import pack_grads, unpack_grads
    gs = [torch.arange(6.0).view(2, 3), None, torch.arange(4.0), torch.ones(1, 2, 2)]
    flat, recipe = pack_grads(gs)
    assert flat.numel() == 6 + 4 + 4
    back = unpack_grads(flat, recipe, len(gs))
    assert back[1] is None
    for a, b in zip(gs, back):
        if a is not None:
            assert a.shape == b.shape and torch.equal(a, b)
    assert pack_grads([None, None])[0] is None


def test_two_rank_row_sharded_aggregation_matches_single_process(tmp_path):
    """Host logic of the row-sharded aggregation (AggregateReduce) over gloo: per-rank CSRs of half of the query
    Gaussians each; the reduced output and gradients equal the unsharded oracle run on both ranks."""
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
    parts = [np.load(os.path.join(str(tmp_path), "agg%d.npz" % r)) for r in range(world)]
    for p in parts:
        assert rel_err(p["out"], out) <= TOL
        for key, wv in zip(AGG_KEYS, want):
            assert rel_err(p["grad_" + key], wv) <= TOL, key
    for key in ["out"] + ["grad_" + k for k in AGG_KEYS]:
        np.testing.assert_array_equal(parts[0][key], parts[1][key])


def test_balanced_ranges_equalise_weight():
    """sharded.balanced_ranges: contiguous windows of equal total weight (host logic, torch on CPU)."""
    import torch

    from pigs_b200.sharded import balanced_ranges, shard_range
    g = torch.Generator().manual_seed(0)
    w = torch.cat([torch.full((4000,), 10), torch.full((6000,), 30), torch.randint(5, 50, (5000,), generator=g)])
    for world in (1, 2, 3, 8):
        rs = balanced_ranges(w, world)
        assert rs[0][0] == 0 and rs[-1][1] == w.numel() and len(rs) == world
        assert all(rs[i][1] == rs[i + 1][0] for i in range(world - 1))
        sums = [int(w[lo:hi].sum()) for lo, hi in rs]
        assert max(sums) - min(sums) <= 2 * int(w.max())          # equal up to one row either side
        even = [int(w[slice(*shard_range(w.numel(), r, world))].sum()) for r in range(world)]
        assert max(sums) <= max(even)
    assert balanced_ranges(torch.zeros(10), 4) == [shard_range(10, r, 4) for r in range(4)]   # no weight: even split
    assert balanced_ranges(torch.zeros(0), 3) == [(0, 0)] * 3
    heavy = torch.tensor([1.0, 1.0, 100.0, 1.0])
    rs = balanced_ranges(heavy, 3)
    assert rs[0][0] == 0 and rs[-1][1] == 4 and all(lo <= hi for lo, hi in rs)


def _local_csr(P, k, seed):
    """A spatially local CSR on P rows: row i names up to k neighbours within +-k of i (ascending), with seeded
    dists / densities — what a kNN list over grid-ordered Gaussians looks like (the C5 bench's 16-regular lists)."""
    rng = np.random.default_rng(seed)
    rows = []
    for i in range(P):
        cand = np.arange(max(0, i - k), min(P, i + k + 1))
        rows.append(np.sort(rng.choice(cand, size=min(k, cand.size), replace=False)))
    cnt = np.array([len(r) for r in rows], np.int64)
    idx = np.concatenate(rows).astype(np.int64)
    M = idx.size
    dists = rng.uniform(-1, 1, (M, 2)).astype(np.float32)
    dens = rng.uniform(0.1, 1.0, M).astype(np.float32)
    return idx, cnt, dists, dens


def _window_csr(idx, cnt, dists, dens, lo, hi):
    starts = np.cumsum(cnt) - cnt
    a0, b0 = int(starts[lo]) if lo < len(cnt) else int(cnt.sum()), int(starts[hi - 1] + cnt[hi - 1]) if hi > lo else 0
    if hi <= lo:
        a0 = b0 = 0
    cnt_r = np.zeros_like(cnt)
    cnt_r[lo:hi] = cnt[lo:hi]
    inv = np.full(len(cnt), np.float32(1e6), np.float32)
    tot = np.add.reduceat(dens, starts)[lo:hi] if hi > lo else np.zeros(0, np.float32)
    inv[lo:hi] = (1.0 / (tot + 1e-6)).astype(np.float32)
    return idx[a0:b0].copy(), np.cumsum(cnt_r).astype(np.int64), dists[a0:b0].copy(), dens[a0:b0].copy(), inv


def _halo_worker(rank, world, port, out_dir, windows, k):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import pigs_oracle as po
        from pigs_b200.sharded import RowHaloPlan, allreduce_grads, pack_grads, unpack_grads
        from tests.fixtures import agg_params
        P = windows[-1][1]
        ap = agg_params(P, 16, 16, 25, 2, seed=5)
        idx, cnt, dists, dens = _local_csr(P, k, seed=6)
        lo, hi = windows[rank]
        csr = _window_csr(idx, cnt, dists, dens, lo, hi)
        plan = RowHaloPlan(P, (lo, hi), torch.from_numpy(csr[0]))
        # full-height staging with NaN outside this rank's window: the exchange must fill exactly what the lists name
        feat = torch.full((P, 16), float("nan"))
        keys = torch.full((P, 16), float("nan"))
        feat[lo:hi] = torch.from_numpy(ap["features"][lo:hi])
        keys[lo:hi] = torch.from_numpy(ap["keys"][lo:hi])
        plan.gather([feat, keys])
        assert torch.isfinite(feat[plan.jlo:plan.jhi]).all() and torch.isfinite(keys[plan.jlo:plan.jhi]).all()
        f_np, k_np = np.nan_to_num(feat.numpy()), np.nan_to_num(keys.numpy())
        a = [f_np, ap["transform"], ap["queries"], k_np, ap["frequencies"], ap["distance_transform"]]
        w, e, f, out = po.aggregate_forward(*a, *csr)
        g = np.random.default_rng(9).normal(size=(P, 16)).astype(np.float32)
        g_loc = np.zeros_like(g)
        g_loc[lo:hi] = g[lo:hi]
        grads = po.aggregate_backward(*a, csr[0], csr[1], csr[2], csr[3], w, e, f, csr[4], g_loc)
        gf = torch.from_numpy(np.ascontiguousarray(grads[0], np.float32))
        gk = torch.from_numpy(np.ascontiguousarray(grads[3], np.float32))
        plan.scatter_add([gf, gk])
        flat, recipe = pack_grads([torch.from_numpy(np.ascontiguousarray(grads[i], np.float32)) for i in (1, 4, 5)])
        allreduce_grads(flat)
        gt, gfr, gdt = unpack_grads(flat, recipe, 3)
        np.savez(os.path.join(out_dir, "halo%d.npz" % rank), lo=lo, hi=hi, dense=int(plan.dense), halo=plan.halo_rows,
                 out=out[lo:hi], g_features=gf[lo:hi].numpy(), g_keys=gk[lo:hi].numpy(), g_queries=grads[2][lo:hi],
                 g_transform=gt.numpy(), g_frequencies=gfr.numpy(), g_distance_transform=gdt.numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
@pytest.mark.parametrize("windows,k", [([(0, 150), (150, 300)], 12),            # halo exchange, two equal windows
                                       ([(0, 90), (90, 200), (200, 330)], 16),  # three unequal windows
                                       ([(0, 30), (30, 60), (60, 90)], 70)])    # strips cover everything: dense fallback
def test_row_halo_plan_matches_single_process(tmp_path, windows, k):
    """Row-parallel aggregation host logic (RowHaloPlan) over gloo: each rank holds only its window of features / keys,
    exchanges the halo rows its lists name, and gets back its own rows of the output and of every gradient; the
    concatenation over ranks equals the unsharded oracle run."""
    from oracle import pigs_oracle as po
    from tests.fixtures import agg_params
    from tests.helpers import TOL, rel_err
    world = len(windows)
    mp.spawn(_halo_worker, args=(world, _free_port(), str(tmp_path), windows, k), nprocs=world, join=True)
    P = windows[-1][1]
    ap = agg_params(P, 16, 16, 25, 2, seed=5)
    idx, cnt, dists, dens = _local_csr(P, k, seed=6)
    csr = _window_csr(idx, cnt, dists, dens, 0, P)
    a = [ap[n] for n in ("features", "transform", "queries", "keys", "frequencies", "distance_transform")]
    w, e, f, out = po.aggregate_forward(*a, *csr)
    g = np.random.default_rng(9).normal(size=(P, 16)).astype(np.float32)
    want = po.aggregate_backward(*a, csr[0], csr[1], csr[2], csr[3], w, e, f, csr[4], g)
    parts = [np.load(os.path.join(str(tmp_path), "halo%d.npz" % r)) for r in range(world)]
    assert all(int(p["dense"]) == int(parts[0]["dense"]) for p in parts)
    assert int(parts[0]["dense"]) == (1 if k >= 40 else 0)
    cat = lambda key: np.concatenate([p[key] for p in parts], 0)
    np.testing.assert_array_equal(cat("out"), out)  # a row's result does not depend on the sharding
    assert rel_err(cat("g_features"), want[0]) <= TOL
    assert rel_err(cat("g_queries"), want[2]) <= TOL
    assert rel_err(cat("g_keys"), want[3]) <= TOL
    for p in parts:
        assert rel_err(p["g_transform"], want[1]) <= TOL
        assert rel_err(p["g_frequencies"], want[4]) <= TOL
        assert rel_err(p["g_distance_transform"], want[5]) <= TOL


def _empty_shard_worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from pigs_b200.sharded import allgather_rows, allreduce_bounds, gather_windows
        D = 2
        if rank == 0:
            mm = torch.tensor([-0.5, -0.25, 0.75, 1.0])
        else:  # a rank without samples joins the collective with the neutral element (ShardedGaussianSampler.preprocess)
            mm = torch.cat([torch.full((D,), float("inf")), torch.full((D,), float("-inf"))])
        allreduce_bounds(mm, D)
        wins = gather_windows((0, 5) if rank == 0 else (5, 5), "cpu")
        buf = torch.zeros(5, 3)
        if rank == 0:
            buf[:] = torch.arange(15.0).view(5, 3)
        allgather_rows(buf, wins)
        np.savez(os.path.join(out_dir, "empty%d.npz" % rank), mm=mm.numpy(), wins=np.array(wins), buf=buf.numpy())
    finally:
        dist.destroy_process_group()


def test_empty_shard_joins_the_collectives(tmp_path):
    mp.spawn(_empty_shard_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    for r in range(2):
        p = np.load(os.path.join(str(tmp_path), "empty%d.npz" % r))
        np.testing.assert_array_equal(p["mm"], np.array([-0.5, -0.25, 0.75, 1.0], np.float32))
        np.testing.assert_array_equal(p["wins"], np.array([[0, 5], [5, 5]]))
        np.testing.assert_array_equal(p["buf"], np.arange(15.0, dtype=np.float32).reshape(5, 3))
