"""Dev tool: neighbour search (preprocess_aggregate) at P Gaussians, C3's recipe — wall time of warm runs and a
checksum of the CSR, so a kernel change can be timed and compared without the full test-suite.

    python tools/nb_breakdown.py [P] [reps]
    ncu --metrics gpu__time_duration.sum --clock-control none -k regex:nb_ --csv --log-file x.csv python tools/nb_breakdown.py 1000000 1
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pigs_b200 import GaussianSampler, _C  # noqa: E402
from tests.fixtures import synthetic  # noqa: E402


def main():
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    torch.cuda.set_device(0)
    rec = synthetic(P, 1024, seed=0)
    d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
    s = GaussianSampler(False)
    s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
    radii = s.radii
    for rep in range(reps):
        ind = rg = ds = dn = it = None
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        ind, rg, ds, dn, it = _C.preprocess_aggregate(d["means"], d["conics"], radii, False)
        b.record()
        torch.cuda.synchronize()
        M = ind.numel()
        # order-sensitive checksum of the lists (position-weighted), plain sums of the float outputs
        w = (torch.arange(M, device="cuda", dtype=torch.int64) % 1000003) + 1
        print("run %d: %.3f ms  M=%d  idx_cs=%d  rng_cs=%d  dists=%.6e dens=%.6e inv=%.6e" % (
            rep, a.elapsed_time(b), M, int(((ind + 2) * w).sum().item() & ((1 << 62) - 1)), int(rg.sum().item()),
            ds.double().abs().sum().item(), dn.double().sum().item(), it.double().sum().item()), flush=True)
        del w


if __name__ == "__main__":
    main()
