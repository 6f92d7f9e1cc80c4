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
