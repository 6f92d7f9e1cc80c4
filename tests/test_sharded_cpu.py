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
