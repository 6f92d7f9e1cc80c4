"""Throughput of the evaluation kernels' other instantiations (1-D, two channels, third derivative, Laplacian output)
next to the headline 2-D one-channel fused kernels.  python tools/bench_variants.py > profiles/rNN_variants.json
Timing: CUDA events, 2 warm-ups, median of 5; pairs = the reference pair set of each case."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pigs_b200 import GaussianSampler, _C  # noqa: E402
from tests.fixtures import synthetic  # noqa: E402


def ev_time(fn, warm=2, it=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(it):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))


# FP32 lane-ops per pair the kernels EXECUTE in their inner loops, counted in the SASS (tools/sass_issue_model.py; FFMA2 /
# FMUL2 / FADD2 = 2): (D, C, mask) -> (forward, backward).  Used as the (conservative) algorithmic count of each variant:
# the roofline fraction below is FP32-pipe utilisation.  DESIGN.md §4 lists them.
EXEC_LANE_OPS = {
    (2, 1, 7): (18.0, 36.5), (2, 1, 19): (16.0, 31.5), (2, 1, 8): (25.1, 52.5), (2, 2, 7): (25.0, 56.0), (2, 2, 8): (30.25, 70.0),
    (2, 1, 1): (10.0, None), (1, 1, 7): (8.0, None),
}
PEAK_T = 148 * 128 * 1.965e9 / 1e12  # T lane-op/s at the maximum SM clock


def case(name, P, N, masks, **kw):
    rec = synthetic(P, N, seed=0, **kw)
    d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
    D, C = rec["means"].shape[1], rec["values"].shape[1]
    s = GaussianSampler(False, fused=False)
    s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
    pairs = s.pair_count()
    nr, bb, sbb, rg, srg = s._sample_state
    a = (d["means"], d["values"], d["conics"], d["samples"], nr)
    rows = []
    for mask in masks:
        def shape(k):
            return (N, C) if (k == 2 and mask & 16) else (N,) + (D,) * k + (C,)
        g = [torch.randn(shape(k), device="cuda") if (mask & (1 << k) or (k == 2 and mask & 16)) else None for k in range(4)]
        tf = ev_time(lambda: _C.sample_forward(mask, *a, bb, sbb, rg, srg, False))
        tb = ev_time(lambda: _C.sample_backward(mask, *a, g, bb, sbb, rg, srg, False))
        row = dict(mask=mask, fwd_ms=tf, bwd_ms=tb, fwd_Gpairs_s=pairs / tf / 1e6, bwd_Gpairs_s=pairs / tb / 1e6)
        ops = EXEC_LANE_OPS.get((D, min(C, 2), mask))
        if ops and C <= 2:
            if ops[0]:
                row["fwd_lane_ops_per_pair"], row["fwd_frac_fp32_roofline"] = ops[0], pairs / (tf * 1e-3) * ops[0] / 1e12 / PEAK_T
            if ops[1]:
                row["bwd_lane_ops_per_pair"], row["bwd_frac_fp32_roofline"] = ops[1], pairs / (tb * 1e-3) * ops[1] / 1e12 / PEAK_T
        rows.append(row)
    return dict(case=name, P=P, N=N, D=D, C=C, pairs=pairs, rows=rows)


def main():
    out = {"gpu": torch.cuda.get_device_name(0),
           "mask_bits": "1 value, 2 gradient, 4 Hessian, 8 third derivative, 16 Laplacian (trace)", "cases": []}
    out["cases"].append(case("2-D C=1 (headline kernels)", 100_000, 1_000_000, (1, 2, 4, 7, 1 | 2 | 16, 8, 15)))
    out["cases"].append(case("2-D C=2", 100_000, 1_000_000, (1, 7, 8), C=2))
    out["cases"].append(case("2-D C=3 (2 + 1 channel groups)", 100_000, 1_000_000, (7,), C=3))
    out["cases"].append(case("1-D C=1", 100_000, 1_000_000, (1, 7, 8), D=1))
    out["cases"].append(case("1-D C=1 at C1's size (1K x 10K)", 1000, 10_000, (7,), D=1))
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
