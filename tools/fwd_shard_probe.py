"""Forward (u + grad u + Hessian) kernel-group time at C3's Gaussians against 4M / 500K / 125K samples (N = 1, a C3 rank at
N = 8, a C5 rank at N = 8) — how much small shards lose to part-filled sample chunks.  PIGS_B200_LIB selects a variant build.
python tools/fwd_shard_probe.py"""
import ctypes
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pigs_b200 import _C, _lib  # noqa: E402
from tests.fixtures import synthetic  # noqa: E402


def main():
    L = _lib.lib()
    rec = synthetic(1_000_000, 4_000_000, seed=0)
    d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
    mm = _C.sample_bounds(d["samples"], 2)
    rows = []
    for n in (4_000_000, 500_000, 125_000):
        smp = d["samples"][:n].contiguous()
        nr, bb, sbb, rg, srg, _ = _C.preprocess_gaussians(d["means"], d["values"], d["covariances"], d["conics"], smp, False, bounds=mm)
        b = rg.numel() // 8 - 1
        g = rg.view(torch.int32)[:2 * b].view(b, 2).to(torch.int64)
        s = srg.view(torch.int32)[:2 * b].view(b, 2).to(torch.int64)
        pairs = int(((g[:, 1] - g[:, 0]) * (s[:, 1] - s[:, 0])).sum())
        a = (d["means"], d["values"], d["conics"], smp, nr)
        for _ in range(2):
            _C.sample_forward(7, *a, bb, sbb, rg, srg, False)
        L.pigs_set_kernel_timing(1)
        L.pigs_kernel_timing_read(0, ctypes.byref(ctypes.c_float()), ctypes.byref(ctypes.c_int()))
        for _ in range(5):
            _C.sample_forward(7, *a, bb, sbb, rg, srg, False)
        torch.cuda.synchronize()
        ms, cnt = ctypes.c_float(), ctypes.c_int()
        L.pigs_kernel_timing_read(0, ctypes.byref(ms), ctypes.byref(cnt))
        L.pigs_set_kernel_timing(0)
        t = ms.value / max(cnt.value, 1)
        rows.append({"samples": n, "pairs": pairs, "fwd_ms": round(t, 4), "Tpairs_per_s": round(pairs / t / 1e9, 4)})
    base = rows[0]["Tpairs_per_s"]
    for r in rows:
        r["rate_vs_full"] = round(r["Tpairs_per_s"] / base, 4)
    print(json.dumps({"lib": os.environ.get("PIGS_B200_LIB", "default"), "rows": rows}))


if __name__ == "__main__":
    main()
