"""Dense all-pairs path (pigs_b200.gaussians.dense_sample -> pigs_dense_forward / _backward) timed on
the GPU next to the reference's pure-PyTorch formulation (oracle/dense_torch.py, the port of gaussians.py:48-116) on the
box's host cores.  C1 = BASELINE.json configs[0]: 1-D, 1K Gaussians x 10K points.  python tools/bench_dense.py"""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
from bench_hbm import timed  # noqa: E402
from oracle import dense_torch as dt  # noqa: E402  (CPU baseline leg only)
from pigs_b200 import gaussians as pg  # noqa: E402


def make(n, N, D, C, seed=0):
    g = torch.Generator().manual_seed(seed)
    means = torch.rand(n, D, generator=g) * 2 - 1
    s = torch.exp(torch.randn(n, D, generator=g) * 0.3 - 4.0)
    if D == 2:
        off = torch.tanh(torch.randn(n, generator=g) * 0.3) * torch.sqrt(s[:, 0] * s[:, 1])
        con = torch.linalg.inv(torch.stack([s[:, 0], off, off, s[:, 1]], -1).reshape(n, 2, 2))
    else:
        con = (1.0 / s).reshape(n, 1, 1)
    ups = [torch.randn(N, C, generator=g), torch.randn(N, D, C, generator=g), torch.randn(N, D, D, C, generator=g)]
    return means, con, torch.rand(n, C, generator=g) * 2 - 1, torch.rand(N, D, generator=g) * 2 - 1, ups


def main():
    torch.cuda.set_device(0)
    rows = []
    for name, n, N, D, C, cpu in (("C1: 1-D 1K x 10K", 1000, 10000, 1, 1, True), ("2-D 1K x 10K", 1000, 10000, 2, 1, True),
                                  ("2-D 20K x 200K", 20000, 200000, 2, 1, False),
                                  ("2-D 100K x 1M (C2's shapes, every pair)", 100000, 1000000, 2, 1, False)):
        means, con, values, samples, ups = make(n, N, D, C)
        m, k, v = (x.cuda().requires_grad_(True) for x in (means, con, values))
        sc, uc = samples.cuda(), [u.cuda() for u in ups]

        def step():
            for t in (m, k, v):
                t.grad = None
            outs = pg.dense_sample(7, m, k, v, sc)
            torch.autograd.backward(outs, uc)

        with torch.no_grad():
            t_f = timed(lambda: pg.dense_sample(7, m, k, v, sc), it=5, warm=2)
        t_s = timed(step, it=5, warm=2)
        row = {"workload": name, "gaussians": n, "points": N, "pairs": n * N, "fwd_ms": t_f, "fwd_bwd_ms": t_s,
               "fwd_bwd_pairs_per_s": n * N / t_s * 1e3}
        if cpu:
            best = 1e30
            for _ in range(5):
                t0 = time.perf_counter()
                dt.step(means, con, values, samples, ups)
                best = min(best, time.perf_counter() - t0)
            row.update({"cpu_port_fwd_bwd_ms": best * 1e3, "cpu_threads": torch.get_num_threads(),
                        "gpu_over_cpu": best * 1e3 / t_s})
        rows.append(row)
    print(json.dumps({"gpu": torch.cuda.get_device_name(0), "functions": "u + grad u + Hessian, one launch each way",
                      "timing": "CUDA events, L2 flushed between repetitions, median of 5", "rows": rows}, indent=1))


if __name__ == "__main__":
    main()
