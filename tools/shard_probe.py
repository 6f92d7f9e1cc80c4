"""Dev tool: the sampling pass of C5 (samples = a contiguous 1/world range of the means, all Gaussians, global tile grid) on\none GPU, to see what a rank of the row-sharded step spends where.  python tools/shard_probe.py"""
import sys, os, json
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tools")
import torch, numpy as np
from bench_hbm import timed
from pigs_b200 import GaussianSampler, _C
from tests.fixtures import synthetic
torch.cuda.set_device(0)
P = 1_000_000
rec = synthetic(P, 16, seed=0)
d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
out = {}
for world in (1, 2, 8):
    n = P // world
    mine = d["means"][:n].contiguous()
    s = GaussianSampler(False)
    mm = _C.sample_bounds(d["means"], 2)  # the global tile grid, as ShardedGaussianSampler all-reduces it
    pre = lambda: s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], mine, bounds=mm)
    t_pre = timed(pre)
    pairs = s.pair_count()
    with torch.no_grad():
        t_f = timed(lambda: s.sample_fused(1 | 2 | 16), it=5, warm=2)
        t_h = timed(lambda: s.sample_fused(7), it=5, warm=2)
    out[world] = dict(samples=n, pairs=pairs, preprocess_ms=t_pre, fwd_lap_ms=t_f, fwd_hess_ms=t_h,
                      Tpairs_per_s_lap=pairs / t_f * 1e-9, ideal_ms_at_full_rate=None)
base = out[1]["pairs"] / out[1]["fwd_lap_ms"]
for w in out:
    out[w]["ideal_ms_at_full_rate"] = out[w]["pairs"] / base
print(json.dumps(out, indent=1))
