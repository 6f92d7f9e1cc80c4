"""Measures the HBM-bound kernels of the path against the copy bandwidth in MEASURED_PEAKS.json:
binning (preprocess), covariance construction, neighbour aggregation forward/backward on a synthetic 16-regular
CSR at P = 1M (config C5 part i) and the neighbour search at reference semantics (C5 part ii, optional: 22 GB CSR).

    python tools/bench_hbm.py [--search] > profiles/rNN_hbm_kernels.json

Timing: CUDA events on torch's current stream (the one the library launches on), 3 warm-ups, L2 flushed between
timed iterations by writing a 512 MB buffer, median of 7.  Algorithmic bytes per unit are the ones DESIGN.md §4 states.
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pigs_b200 import GaussianSampler, _C  # noqa: E402
from pigs_b200 import gaussians as pg  # noqa: E402
from tests.fixtures import agg_params, synthetic  # noqa: E402

_flush = None


def timed(fn, it=7, warm=3):
    global _flush
    if _flush is None:
        _flush = torch.empty(128 * 1024 * 1024, dtype=torch.float32, device="cuda")
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(it):
        _flush.fill_(1.0)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))


def peak_gbs():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    except Exception:
        return 6550.0, "fallback: B200_PROFILING.md measured copy bandwidth"


def row(name, ms, nbytes, what):
    pk, _ = peak_gbs()
    gbs = nbytes / (ms * 1e-3) / 1e9
    return {"kernel": name, "ms": round(ms, 4), "algorithmic_MB": round(nbytes / 1e6, 1), "achieved_GBs": round(gbs, 1),
            "frac_of_hbm_peak": round(gbs / pk, 3), "bytes_model": what}


def regular_csr(means, k):
    """k nearest Gaussians of every Gaussian incl. itself (Euclidean, no wrap), ascending index per row."""
    from scipy.spatial import cKDTree
    _, idx = cKDTree(means).query(means, k=k, workers=-1)
    idx = np.sort(idx.astype(np.int64), axis=1)
    return idx.reshape(-1), np.arange(1, means.shape[0] + 1, dtype=np.int64) * k


def main():
    torch.cuda.set_device(0)
    out = {"gpu": torch.cuda.get_device_name(0), "peak_GBs": peak_gbs()[0], "peak_source": peak_gbs()[1], "rows": []}
    only = [a.split("=", 1)[1] for a in sys.argv if a.startswith("--only=")]
    only = only[0] if only else None  # --only=agg | --only=binning: just those rows (variant sweeps)
    P, N = 1_000_000, 4_000_000
    rec = synthetic(P, N if only != "agg" else 16, seed=0)
    d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
    s = GaussianSampler(False)

    # ---- binning (C3): S1-S7 ----
    ms = timed(lambda: s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"]))
    R = s.num_rendered
    nbytes = 28 * P + (8 + 16) * R + 4 * R + (8 + 8) * N + (4 + 16) * N + 4 * N
    out["rows"].append(row("preprocess (tile binning + per-preprocess class staging, 14 launches, one D2H header readback)", ms, nbytes,
                           "28 B/Gaussian + 28 B/instance (emit 8, hist 4, scatter 16) + 40 B/sample (key 16, hist 4, scatter 16, ranges 4); "
                           "P=%d R=%d N=%d" % (P, R, N)))

    if only == "binning":
        print(json.dumps(out, indent=1))
        return
    if only != "agg":
        rows_other(out, P, N)
    rows_agg(out, rec, d, s, P)
    print(json.dumps(out, indent=1))


def rows_other(out, P, N):
    # ---- covariance construction ----
    sc = torch.rand(P, 2, device="cuda") * 1e-3 + 1e-4
    tr = torch.randn(P, 1, device="cuda") * 0.3
    ms = timed(lambda: _C.build_covariances(sc, tr))
    out["rows"].append(row("cov_forward_kernel", ms, 36 * P, "12 B read + 24 B written per Gaussian (output allocation included in the time)"))
    gc, gk = torch.randn(P, 3, device="cuda"), torch.randn(P, 3, device="cuda")
    ms = timed(lambda: _C.build_covariances_backward(sc, tr, gc, gk))
    out["rows"].append(row("cov_backward_kernel", ms, 48 * P, "36 B read + 12 B written per Gaussian"))

    # ---- PDE residual loss kernels at N = 4M samples (model_pn.py:612-631, 822-845) ----
    g1 = torch.ones((), device="cuda")
    u1, ux1, lap1, ut1 = (torch.randn(N, 1, device="cuda"), torch.randn(N, 2, 1, device="cuda"),
                          torch.randn(N, 1, device="cuda"), torch.randn(N, 1, device="cuda"))
    ms = timed(lambda: _C.pde_residual_forward(_C.PDE_BURGERS, u1, ux1, lap1, ut1, 0.01, 0.01))
    out["rows"].append(row("residual_forward_kernel + finish (BURGERS, C=1)", ms, 16 * N, "16 B read per sample (u, ux0, lap, ut)"))
    ms = timed(lambda: _C.pde_residual_backward(_C.PDE_BURGERS, u1, ux1, lap1, ut1, 0.01, 0.01, g1))
    out["rows"].append(row("residual_backward_kernel (BURGERS, C=1)", ms, 36 * N, "16 B read + 20 B written per sample (d_u, d_ux[2], d_lap, d_ut)"))
    u2, ux2, wx2, wxx2, wt2 = (torch.randn(N, 2, device="cuda"), torch.randn(N, 2, 2, device="cuda"),
                               torch.randn(N, 2, device="cuda"), torch.randn(N, 2, 2, device="cuda"), torch.randn(N, device="cuda"))
    ms = timed(lambda: _C.pde_ns_residual_forward(u2, ux2, wx2, wxx2, wt2, 0.01, 0.01))
    out["rows"].append(row("ns_residual_forward_kernel + finish (NAVIER_STOKES)", ms, 52 * N, "52 B read per sample (u 8, ux 16, wx 8, wxx 16, wt 4)"))
    ms = timed(lambda: _C.pde_ns_residual_backward(u2, ux2, wx2, wxx2, wt2, 0.01, 0.01, g1))
    out["rows"].append(row("ns_residual_backward_kernel (NAVIER_STOKES)", ms, 104 * N, "52 B read + 52 B written per sample"))
    uxxx2 = torch.randn(N, 2, 2, 2, 2, device="cuda")
    ms = timed(lambda: _C.vorticity_forward(uxxx2))
    out["rows"].append(row("vorticity_forward_kernel (uxxx -> wxx)", ms, 80 * N, "4 groups per sample: 16 B read + 4 B written each"))
    del u1, ux1, lap1, ut1, u2, ux2, wx2, wxx2, wt2, uxxx2



def rows_agg(out, rec, d, s, P):
    # ---- aggregation on a 16-regular CSR (C5 i): L = K = 16, E = 25, nF = 6 ----
    k = 16
    idx, rng = regular_csr(rec["means"], k)
    indices, ranges = torch.from_numpy(idx).cuda(), torch.from_numpy(rng).cuda()
    M = indices.numel()
    radii = s.radii
    i_of = torch.arange(P, device="cuda").repeat_interleave(k)
    X = d["means"][indices] - d["means"][i_of]
    X = torch.where(X.abs() > 1, X - 2 * torch.sign(X), X)
    dists = (X / (0.333 * radii[i_of] + 1e-6).unsqueeze(1)).contiguous()
    cj = d["conics"][indices]
    power = -0.5 * (cj[:, 0] * X[:, 0] ** 2 + cj[:, 2] * X[:, 1] ** 2) - cj[:, 1] * X[:, 0] * X[:, 1]
    dens = torch.exp(power).contiguous()
    inv_total = (1.0 / (dens.view(P, k).sum(1) + 1e-6)).contiguous()
    ap = {kk: torch.from_numpy(v).cuda() for kk, v in agg_params(P, 16, 16, 25, 2, seed=100).items()}
    a = (ap["features"], ap["transform"], ap["queries"], ap["keys"], ap["frequencies"], ap["distance_transform"])
    ms = timed(lambda: _C.aggregate_neighbors(*a, indices, ranges, dists, dens, inv_total, False))
    out["rows"].append(row("agg_forward_kernel (outputs allocated, not filled)", ms, 160 * M + 140 * P,
                           "160 B/slot (idx 8, dist 8, density 4, feature row 64, key row 64, 3 saved 12) + 140 B/Gaussian; M=%d" % M))
    w, e, f, o = _C.aggregate_neighbors(*a, indices, ranges, dists, dens, inv_total, False)
    g = torch.randn_like(o)
    ms = timed(lambda: _C.aggregate_neighbors_backward(*a, indices, ranges, dists, dens, w, e, f, inv_total, g, False))
    out["rows"].append(row("agg_backward_kernel (+ zero fills of its 6 outputs)", ms, 416 * M + 260 * P,
                           "416 B/slot (fwd reads 160 + atomic RMW of dL/dfeature and dL/dkey rows 256) + 260 B/Gaussian"))

    # ---- neighbour search at reference semantics (C5 ii) ----
    if "--search" in sys.argv:
        times = []
        for rep in range(3):  # the first run pays cudaMalloc of the ~32 GB CSR; the caching allocator serves the next ones
            ind = rg = ds = dn = it = None
            torch.cuda.synchronize()
            a0, b0 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record()
            ind, rg, ds, dn, it = _C.preprocess_aggregate(d["means"], d["conics"], radii, False)
            b0.record()
            torch.cuda.synchronize()
            times.append(a0.elapsed_time(b0))
        ms = min(times[1:])
        out["neighbour_search_runs_ms"] = [round(t, 3) for t in times]
        Ms = ind.numel()
        out["rows"].append(row("preprocess_aggregate (grid neighbour search + CSR write-out; best warm run of 2)", ms, 20 * Ms,
                               "CSR write-out 20 B/slot (idx 8, dist 8, density 4); M=%d slots = %.0f per Gaussian" % (Ms, Ms / P)))


if __name__ == "__main__":
    main()
