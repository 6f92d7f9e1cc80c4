"""TEST INFRASTRUCTURE ONLY — ctypes front-end of the CPU oracle (oracle/pigs_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; the product (pigs_b200/) never does.  numpy in, numpy out.
Each function names the reference entry point it restates (paths relative to
/root/reference/diff-gaussian-sampling).
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

FN_VALUE, FN_DERIV, FN_HESS, FN_THIRD = 0, 1, 2, 3


def build():
    """Compile oracle/pigs_oracle.c -> oracle/libpigs_oracle.so (gcc, OpenMP)."""
    subprocess.run(["make", "-s", "-C", _HERE, "libpigs_oracle.so"], check=True)


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "libpigs_oracle.so")
        if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(os.path.join(_HERE, "pigs_oracle.c")):
            build()
        _LIB = ctypes.CDLL(so)
        _LIB.pigs_oracle_tile_grid.restype = ctypes.c_int
        _LIB.pigs_oracle_bin_gaussians.restype = ctypes.c_int64
        _LIB.pigs_oracle_neighbor_counts.restype = ctypes.c_int64
    return _LIB


def _f32(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float32))


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


_i64 = ctypes.c_int64
_int = ctypes.c_int


def out_shape(fn, N, D, C):
    return (N,) + (D,) * fn + (C,)


def preprocess(means, covariances, samples, recip_mode=1, bounds=None, cull=None):
    """PreprocessCUDA (sample_points.cu:37-98) + Sampler::preprocess (sampler_impl.cu:216-330).
    cull=(sigmas, k) (NOT reference semantics): the product's opt-in culled binning — the reference pair set
    intersected with a per-dimension `sigmas`-sigma box test on k x k sub-tiles of every reference tile; the returned
    grid / blocks / lists / ranges are then the fine ones, `radii` stays the reference's (see pigs_oracle.c).

    Returns a dict with min_b, max_b, grid, blocks, radii, touched, num_rendered, point_list,
    ranges [(blocks+1),2] u32, sample_list, sample_tiles, sample_ranges.  `bounds=(min_b,max_b)`
    overrides the per-call sample bounding box (multi-GPU: global bounds)."""
    L = lib()
    means = _f32(means)
    P = means.shape[0]
    D = means.shape[-1] if means.ndim > 1 else 1
    means = means.reshape(P, D)
    cov = _f32(covariances).reshape(P, D * (D + 1) // 2)
    samples = _f32(samples).reshape(-1, D)
    N = samples.shape[0]
    min_b = np.zeros(D, np.float32)
    max_b = np.zeros(D, np.float32)
    if bounds is None:
        L.pigs_oracle_bounds(_p(samples), _i64(N), _int(D), _p(min_b), _p(max_b))
    else:
        min_b[:] = bounds[0]
        max_b[:] = bounds[1]
    grid = np.zeros(D, np.int32)
    blocks = L.pigs_oracle_tile_grid(_p(min_b), _p(max_b), _int(D), _int(recip_mode), _p(grid))
    radii = np.zeros(P, np.float32)
    touched = np.zeros(P, np.uint32)
    L.pigs_oracle_radii(_i64(P), _int(D), _p(means), _p(cov), _p(min_b), _p(grid), _p(radii), _p(touched))
    if cull is not None:
        return _preprocess_culled(L, cull, means, cov, samples, P, D, N, min_b, max_b, grid, blocks, radii)
    R = int(touched.astype(np.int64).sum())
    point_list = np.zeros(max(R, 1), np.uint32)
    ranges = np.zeros((blocks + 1, 2), np.uint32)
    R2 = L.pigs_oracle_bin_gaussians(_i64(P), _int(D), _p(means), _p(radii), _p(min_b), _p(grid), _int(blocks),
                                     _p(point_list), _p(ranges))
    if R2 < 0:
        raise ValueError("input hits the reference's out-of-range tile alias (sampler_impl.cu:108-110)")
    assert R2 == R, (R2, R)
    sample_list = np.zeros(max(N, 1), np.uint32)
    sample_tiles = np.zeros(max(N, 1), np.uint32)
    sample_ranges = np.zeros((blocks + 1, 2), np.uint32)
    L.pigs_oracle_bin_samples(_i64(N), _int(D), _p(samples), _p(min_b), _p(grid), _int(blocks), _p(sample_list),
                              _p(sample_tiles), _p(sample_ranges))
    return dict(P=P, D=D, N=N, min_b=min_b, max_b=max_b, grid=grid, blocks=blocks, radii=radii, touched=touched,
                num_rendered=R, point_list=point_list[:R], ranges=ranges, sample_list=sample_list[:N],
                sample_tiles=sample_tiles[:N], sample_ranges=sample_ranges)


def _preprocess_culled(L, cull, means, cov, samples, P, D, N, min_b, max_b, ref_grid, ref_blocks, radii):
    sig, k = float(cull[0]), int(cull[1])
    L.pigs_oracle_bin_gaussians_culled.restype = ctypes.c_int64
    args = (_i64(P), _int(D), _p(means), _p(cov), _p(radii), _p(min_b), _p(ref_grid), _int(k), ctypes.c_float(sig))
    R = L.pigs_oracle_bin_gaussians_culled(*args, None, None)
    blocks = ref_blocks * (k ** D)
    point_list = np.zeros(max(R, 1), np.uint32)
    ranges = np.zeros((blocks + 1, 2), np.uint32)
    R2 = L.pigs_oracle_bin_gaussians_culled(*args, _p(point_list), _p(ranges))
    assert R2 == R
    sample_list = np.zeros(max(N, 1), np.uint32)
    sample_tiles = np.zeros(max(N, 1), np.uint32)
    sample_ranges = np.zeros((blocks + 1, 2), np.uint32)
    L.pigs_oracle_bin_samples_culled(_i64(N), _int(D), _p(samples), _p(min_b), _p(ref_grid), _int(k), _p(sample_list),
                                     _p(sample_tiles), _p(sample_ranges))
    return dict(P=P, D=D, N=N, min_b=min_b, max_b=max_b, grid=(ref_grid * k).astype(np.int32), ref_grid=ref_grid,
                blocks=blocks, radii=radii, num_rendered=int(R), point_list=point_list[:R], ranges=ranges,
                sample_list=sample_list[:N], sample_tiles=sample_tiles[:N], sample_ranges=sample_ranges)


def pair_count(st):
    """Size of the reference pair set: sum over tiles of n_samples(tile) * n_gaussians(tile)."""
    b = st["blocks"]
    g = st["ranges"][:b].astype(np.int64)
    s = st["sample_ranges"][:b].astype(np.int64)
    return int(((g[:, 1] - g[:, 0]) * (s[:, 1] - s[:, 0])).sum())


def forward(fn, st, means, values, conics, samples, acc64=False):
    """SampleGaussians{,Derivative,Laplacian,Third}CUDA (sample_points.cu:198-297) -> renderCUDA<F>
    (cuda_sampler/forward.cu:88-275).  Output [N, D^fn.., C] fp32."""
    L = lib()
    P, D, N = st["P"], st["D"], st["N"]
    means = _f32(means).reshape(P, D)
    values = _f32(values).reshape(P, -1)
    C = values.shape[1]
    conics = _f32(conics).reshape(P, D * (D + 1) // 2)
    samples = _f32(samples).reshape(N, D)
    out = np.zeros(out_shape(fn, N, D, C), np.float32)
    if P and N:
        L.pigs_oracle_forward(_int(fn), _int(D), _int(C), _int(st["blocks"]), _p(st["ranges"]),
                              _p(st["sample_ranges"]), _p(np.ascontiguousarray(st["point_list"])),
                              _p(np.ascontiguousarray(st["sample_list"])), _p(means), _p(values), _p(conics),
                              _p(samples), _p(out), _int(1 if acc64 else 0))
    return out


def backward(fn, st, means, values, conics, samples, dL_dout):
    """SampleGaussians*BackwardCUDA (sample_points.cu:146-196,299-372) -> backward renderCUDA<F>
    (cuda_sampler/backward.cu:27-416).  Returns dL_dmeans [P,D], dL_dvalues [P,C], dL_dconics [P,D(D+1)/2]."""
    L = lib()
    P, D, N = st["P"], st["D"], st["N"]
    means = _f32(means).reshape(P, D)
    values = _f32(values).reshape(P, -1)
    C = values.shape[1]
    K = D * (D + 1) // 2
    conics = _f32(conics).reshape(P, K)
    samples = _f32(samples).reshape(N, D)
    g = _f32(dL_dout).reshape(out_shape(fn, N, D, C))
    dm = np.zeros((P, D), np.float32)
    dv = np.zeros((P, C), np.float32)
    dc = np.zeros((P, K), np.float32)
    if P and N:
        L.pigs_oracle_backward(_int(fn), _int(P), _int(D), _int(C), _int(st["blocks"]), _p(st["ranges"]),
                               _p(st["sample_ranges"]), _p(np.ascontiguousarray(st["point_list"])),
                               _p(np.ascontiguousarray(st["sample_list"])), _p(means), _p(values), _p(conics),
                               _p(samples), _p(g), _p(dm), _p(dv), _p(dc))
    return dm, dv, dc


def dense(which, means, conics_full, values, samples):
    """/root/reference/gaussians.py:48-58, 89-101, 103-116 (dense, no cutoff, no torus, full conics).
    which: bit0 value, bit1 derivative, bit2 second derivative. Returns (u, du, ddu)."""
    L = lib()
    means = _f32(means)
    n, d = means.shape
    conics_full = _f32(conics_full).reshape(n, d, d)
    values = _f32(values).reshape(n, -1)
    C = values.shape[1]
    samples = _f32(samples).reshape(-1, d)
    N = samples.shape[0]
    u = np.zeros((N, C), np.float32)
    du = np.zeros((N, d, C), np.float32)
    ddu = np.zeros((N, d, d, C), np.float32)
    L.pigs_oracle_dense(_int(which), _i64(n), _int(d), _int(C), _i64(N), _p(means), _p(conics_full), _p(values),
                        _p(samples), _p(u), _p(du), _p(ddu))
    return u, du, ddu


def neighbors(means, conics, radii):
    """AggregateNeighborsPreprocessCUDA (aggregate_neighbors.cu:323-367): findCollisions + cumsum + preprocess.
    Returns indices i64[M], ranges i64[P] (inclusive cumsum), dists f32[M,D], densities f32[M], inv_total f32[P]."""
    L = lib()
    means = _f32(means)
    P, D = means.shape
    conics = _f32(conics).reshape(P, D * (D + 1) // 2)
    radii = _f32(radii).reshape(P)
    counts = np.zeros(P, np.int64)
    M = L.pigs_oracle_neighbor_counts(_i64(P), _int(D), _p(means), _p(radii), _p(counts))
    ranges = np.cumsum(counts).astype(np.int64)
    indices = np.full(M, -1, np.int64)
    dists = np.zeros((M, D), np.float32)
    densities = np.zeros(M, np.float32)
    inv_total = np.zeros(P, np.float32)
    L.pigs_oracle_neighbor_fill(_i64(P), _int(D), _p(means), _p(conics), _p(radii), _p(ranges), _p(indices),
                                _p(dists), _p(densities), _p(inv_total))
    return indices, ranges, dists, densities, inv_total


def neighbor_rows(means, conics, radii, rows, cap=8192):
    """`neighbors` for a sample of rows (each still scans all P candidates).  Returns counts i64[n], indices i64[n,cap]
    (-1 beyond the count), dists f32[n,cap,D], densities f32[n,cap], inv_total f32[n]."""
    L = lib()
    means = _f32(means)
    P, D = means.shape
    conics = _f32(conics).reshape(P, D * (D + 1) // 2)
    radii = _f32(radii).reshape(P)
    rows = np.ascontiguousarray(rows, np.int64)
    n = rows.shape[0]
    counts = np.zeros(n, np.int64)
    indices = np.full((n, cap), -1, np.int64)
    dists = np.zeros((n, cap, D), np.float32)
    densities = np.zeros((n, cap), np.float32)
    inv_total = np.zeros(n, np.float32)
    L.pigs_oracle_neighbor_rows(_i64(P), _int(D), _p(means), _p(conics), _p(radii), _i64(n), _p(rows), _i64(cap),
                                _p(counts), _p(indices), _p(dists), _p(densities), _p(inv_total))
    return counts, indices, dists, densities, inv_total


def aggregate_forward(features, transform, queries, keys, frequencies, distance_transform, indices, ranges, dists,
                      densities, inv_total):
    """AggregateNeighborsCUDA (aggregate_neighbors.cu:369-415) -> aggregateNeighbors (:129-207).
    Returns weights, embeddings, factors [M], neighbor_features [P,L]."""
    L_ = lib()
    features = _f32(features)
    P, L = features.shape
    transform = _f32(transform)
    queries = _f32(queries)
    keys = _f32(keys)
    K = queries.shape[-1]
    frequencies = _f32(frequencies)
    dt = _f32(distance_transform)
    E = dt.shape[-1] // 2
    dists = _f32(dists)
    D = dists.shape[-1]
    M = densities.shape[0]
    indices = np.ascontiguousarray(indices, np.int64)
    ranges = np.ascontiguousarray(ranges, np.int64)
    densities = _f32(densities)
    inv_total = _f32(inv_total)
    w = np.zeros(M, np.float32)
    e = np.zeros(M, np.float32)
    f = np.zeros(M, np.float32)
    out = np.zeros((P, L), np.float32)
    L_.pigs_oracle_aggregate_forward(_i64(P), _int(D), _int(L), _int(K), _int(E), _p(features), _p(transform),
                                     _p(queries), _p(keys), _p(frequencies), _p(dt), _p(indices), _p(ranges),
                                     _p(dists), _p(densities), _p(inv_total), _p(w), _p(e), _p(f), _p(out))
    return w, e, f, out


def aggregate_backward(features, transform, queries, keys, frequencies, distance_transform, indices, ranges, dists,
                       densities, weights, embeddings, factors, inv_total, dL_dout):
    """AggregateNeighborsBackwardCUDA (aggregate_neighbors.cu:417-475) -> aggregateNeighborsBackward (:210-321)."""
    L_ = lib()
    features = _f32(features)
    P, L = features.shape
    transform = _f32(transform)
    queries = _f32(queries)
    keys = _f32(keys)
    K = queries.shape[-1]
    frequencies = _f32(frequencies)
    dt = _f32(distance_transform)
    E = dt.shape[-1] // 2
    dists = _f32(dists)
    D = dists.shape[-1]
    indices = np.ascontiguousarray(indices, np.int64)
    ranges = np.ascontiguousarray(ranges, np.int64)
    gf = np.zeros_like(features)
    gt = np.zeros_like(transform)
    gq = np.zeros_like(queries)
    gk = np.zeros_like(keys)
    gfr = np.zeros_like(frequencies)
    gdt = np.zeros_like(dt)
    L_.pigs_oracle_aggregate_backward(_i64(P), _int(D), _int(L), _int(K), _int(E), _p(features), _p(transform),
                                      _p(queries), _p(keys), _p(frequencies), _p(dt), _p(indices), _p(ranges),
                                      _p(dists), _p(_f32(densities)), _p(_f32(weights)), _p(_f32(embeddings)),
                                      _p(_f32(factors)), _p(_f32(inv_total)), _p(_f32(dL_dout)), _p(gf), _p(gt),
                                      _p(gq), _p(gk), _p(gfr), _p(gdt))
    return gf, gt, gq, gk, gfr, gdt


# ---------------------------------------------------------------------------------------------------------
# covariance / conic construction — numpy float64 restatement of gaussians.py:163-193
# ---------------------------------------------------------------------------------------------------------
def build_covariances(s, t):
    """gaussians.py:164 tau = tanh(t) * sqrt(prod(s)); :165-177 Sigma = [[s0, tau], [tau, s1]];
    :181 conic = inverse(Sigma); :186-189 flatten [0, 1, 3].  Returns float64 (cov, conic), [P, 3] (2-D) or [P, 1]."""
    s = np.asarray(s, np.float64)
    if s.shape[-1] == 1:
        return s.copy(), 1.0 / s
    t = np.asarray(t, np.float64).reshape(-1)
    tau = np.tanh(t) * np.sqrt(s[:, 0] * s[:, 1])
    det = s[:, 0] * s[:, 1] - tau * tau
    cov = np.stack([s[:, 0], tau, s[:, 1]], 1)
    con = np.stack([s[:, 1] / det, -tau / det, s[:, 0] / det], 1)
    return cov, con


def build_covariances_backward(s, t, g_cov, g_con):
    """Gradient of the chain above as torch.autograd computes it for the reference: dL/dSigma = -K^T G K^T with the
    flattened conic gradient placed at entries [0,0], [0,1], [1,1] (gaussians.py:181,186-189), both off-diagonal
    entries of Sigma being the same tau (:176-177).  float64; returns (dL/ds [P, D], dL/dt [P, 1])."""
    s = np.asarray(s, np.float64)
    g_cov = np.asarray(g_cov, np.float64)
    g_con = np.asarray(g_con, np.float64)
    if s.shape[-1] == 1:
        return g_cov - g_con / (s * s), np.zeros((s.shape[0], 1))
    t = np.asarray(t, np.float64).reshape(-1)
    P = s.shape[0]
    u = np.tanh(t)
    r = np.sqrt(s[:, 0] * s[:, 1])
    tau = u * r
    Sig = np.empty((P, 2, 2))
    Sig[:, 0, 0], Sig[:, 1, 1] = s[:, 0], s[:, 1]
    Sig[:, 0, 1] = Sig[:, 1, 0] = tau
    K = np.linalg.inv(Sig)
    G = np.zeros((P, 2, 2))
    G[:, 0, 0], G[:, 0, 1], G[:, 1, 1] = g_con[:, 0], g_con[:, 1], g_con[:, 2]
    KT = np.transpose(K, (0, 2, 1))
    dS = -KT @ G @ KT
    dS[:, 0, 0] += g_cov[:, 0]
    dS[:, 0, 1] += g_cov[:, 1]
    dS[:, 1, 1] += g_cov[:, 2]
    dtau = dS[:, 0, 1] + dS[:, 1, 0]
    ds0 = dS[:, 0, 0] + dtau * u * s[:, 1] / (2 * r)
    ds1 = dS[:, 1, 1] + dtau * u * s[:, 0] / (2 * r)
    return np.stack([ds0, ds1], 1), (dtau * (1 - u * u) * r).reshape(-1, 1)


# ---- Laplacian + PDE residual -------------------------------------------------
PDE_DIFFUSION, PDE_BURGERS, PDE_WAVE = 1, 3, 4  # model_pn.py:13-19 Problem enum values


def laplacian(st, means, values, conics, samples):
    """`uxx[:,0,0] + uxx[:,1,1]` of the reference's Hessian output (model_pn.py:613-614), [N, C]."""
    H = forward(FN_HESS, st, means, values, conics, samples)
    D = st["D"]
    out = H[:, 0, 0, :].copy() if D == 2 else H.reshape(H.shape[0], -1).copy()
    if D == 2:
        out = out + H[:, 1, 1, :]
    return out.astype(np.float32)


def laplacian_backward(st, means, values, conics, samples, dL_dlap):
    """autograd of the trace: the upstream gradient goes on the diagonal of the Hessian's, zeros elsewhere."""
    D, N = st["D"], st["N"]
    g = _f32(dL_dlap).reshape(N, -1)
    C = g.shape[1]
    G = np.zeros(out_shape(FN_HESS, N, D, C), np.float32)
    for d in range(D):
        G[(slice(None),) + (d, d)] = g
    return backward(FN_HESS, st, means, values, conics, samples, G)


def pde_rhs(problem, u, ux, lap, nu):
    """model_pn.py:612-627 pde_rhs with the trace already taken; fp32 op by op like the torch expressions."""
    f = np.float32
    if problem == PDE_DIFFUSION:
        return lap.astype(f)
    if problem == PDE_BURGERS:
        return (f(nu) * lap - u * ux[:, 0]).astype(f)
    if problem == PDE_WAVE:
        return np.stack((u[..., 1], f(10) * lap[..., 0] - f(0.1) * u[..., 1]), axis=-1).astype(f)
    raise ValueError("Unexpected PDE problem: %r" % (problem,))


def pde_residual_loss(problem, u, ux, lap, ut, nu, dt):
    """model_pn.py:822-837: rhs = dt * pde_rhs; loss = mean((ut - rhs)**2) (WAVE: 0.01 * mean + mean).
    Returns (residual [N, C] fp32, loss float64-accumulated)."""
    r = (ut - np.float32(dt) * pde_rhs(problem, u, ux, lap, nu)).astype(np.float32)
    sq = r.astype(np.float64) ** 2
    if problem == PDE_WAVE:
        return r, 0.01 * sq[..., 0].mean() + sq[..., 1].mean()
    return r, sq.mean()


def pde_residual_backward(problem, u, ux, lap, ut, nu, dt, g_loss):
    """Analytic gradient of pde_residual_loss (checked against torch.autograd through the real model_pn.pde_rhs by
    tests/golden/pde_golden.npz).  Returns (d_u, d_ux, d_lap, d_ut), float64."""
    r, _ = pde_residual_loss(problem, u, ux, lap, ut, nu, dt)
    r = r.astype(np.float64)
    N = r.shape[0]
    z = np.zeros_like
    if problem == PDE_WAVE:
        dr = np.stack((0.01 * 2 * r[..., 0] / N, 2 * r[..., 1] / N), -1) * g_loss
        d_u, d_lap = z(dr), z(dr)
        d_u[..., 1] = -dt * dr[..., 0] + 0.1 * dt * dr[..., 1]
        d_lap[..., 0] = -10.0 * dt * dr[..., 1]
        return d_u, (None if ux is None else np.zeros(ux.shape)), d_lap, dr
    dr = 2 * r / r.size * g_loss
    drhs = -dt * dr
    if problem == PDE_DIFFUSION:
        return z(dr), (None if ux is None else np.zeros(ux.shape)), drhs, dr
    d_ux = np.zeros(ux.shape)
    d_ux[:, 0] = -u * drhs
    return -ux[:, 0] * drhs, d_ux, nu * drhs, dr


def vorticity(x):
    """model_pn.py:779-781: `ux[:,0,1] - ux[:,1,0]`, `uxx[...,0,1] - uxx[...,1,0]`, `uxxx[...,0,1] - uxxx[...,1,0]` — the
    trailing two axes of a C = 2 sampler output are (last spatial axis, channel)."""
    x = _f32(x)
    return (x[..., 0, 1] - x[..., 1, 0]).astype(np.float32)


def vorticity_backward(g):
    g = np.asarray(g, np.float64)
    d = np.zeros(g.shape + (2, 2))
    d[..., 0, 1] = g
    d[..., 1, 0] = -g
    return d


def ns_residual_loss(u, ux, wx, wxx, wt, nu, dt):
    """NAVIER_STOKES, vorticity form.  model_pn.py:629-631 `nu*(wxx00 + wxx11) - (u0*wx0 + u1*wx1)`, :824 `rhs = dt*...`,
    :843-845 `mean((ux00 + ux11)**2) + mean((wt - rhs)**2)`; fp32 op by op like the torch expressions.
    Returns (residual [N] fp32, divergence [N] fp32, loss float64-accumulated)."""
    f = np.float32
    u, ux, wx, wxx, wt = (_f32(t) for t in (u, ux, wx, wxx, wt))
    rhs = (f(nu) * (wxx[:, 0, 0] + wxx[:, 1, 1]) - (u[:, 0] * wx[:, 0] + u[:, 1] * wx[:, 1])).astype(f)
    r = (wt.reshape(-1) - f(dt) * rhs).astype(f)
    div = (ux[:, 0, 0] + ux[:, 1, 1]).astype(f)
    return r, div, (div.astype(np.float64) ** 2).mean() + (r.astype(np.float64) ** 2).mean()


def ns_residual_backward(u, ux, wx, wxx, wt, nu, dt, g_loss):
    """Analytic gradient of ns_residual_loss (checked against torch.autograd through the real model_pn.pde_rhs by
    tests/golden/pde_golden.npz).  Returns (d_u, d_ux, d_wx, d_wxx, d_wt), float64."""
    r, div, _ = ns_residual_loss(u, ux, wx, wxx, wt, nu, dt)
    N = r.shape[0]
    dr = 2.0 * r.astype(np.float64) / N * g_loss
    dd = 2.0 * div.astype(np.float64) / N * g_loss
    drhs = -float(np.float32(dt)) * dr
    u, wx = np.asarray(u, np.float64).reshape(N, 2), np.asarray(wx, np.float64).reshape(N, 2)
    d_ux, d_wxx = np.zeros((N, 2, 2)), np.zeros((N, 2, 2))
    d_ux[:, 0, 0] = d_ux[:, 1, 1] = dd
    d_wxx[:, 0, 0] = d_wxx[:, 1, 1] = float(np.float32(nu)) * drhs
    return -wx * drhs[:, None], d_ux, -u * drhs[:, None], d_wxx, dr


def time_blend(ts, x1, x0):
    """model_pn.py:795-805 TRAPEZOID `ts.reshape(-1,1,..) * x[-1] + (1 - ts.reshape(-1,1,..)) * x[-2]`, fp32 op by op;
    a python number stands for the FORWARD (0 -> x[-2]) / BACKWARD (1 -> x[-1]) rules (:806-813)."""
    x1, x0 = _f32(x1), _f32(x0)
    if np.isscalar(ts):
        return x1.copy() if float(ts) == 1.0 else x0.copy()
    t = _f32(ts).reshape((-1,) + (1,) * (x1.ndim - 1))
    return (t * x1 + (np.float32(1) - t) * x0).astype(np.float32)


def _blend_adjoint(ts, g, ndim):
    if np.isscalar(ts):
        return (g, np.zeros_like(g)) if float(ts) == 1.0 else (np.zeros_like(g), g)
    t = np.asarray(ts, np.float32).astype(np.float64).reshape((-1,) + (1,) * (ndim - 1))
    return t * g, (np.float32(1) - np.asarray(ts, np.float32)).astype(np.float64).reshape(t.shape) * g


def pde_residual_loss_timelevels(problem, ts, u, ux, lap, nu, dt):
    """compute_loss with two time levels (model_pn.py:794-837): blended u / ux / lap into pde_rhs, ut = u[-1] - u[-2].
    u, ux, lap: pairs (level [-1], level [-2]).  (The reference blends the Hessians and then takes the trace; blending
    the traces differs by rounding only.)"""
    ut = (_f32(u[0]) - _f32(u[1])).astype(np.float32)
    us = time_blend(ts, *u)
    uxs = None if ux is None else time_blend(ts, *ux)
    return pde_residual_loss(problem, us, uxs, time_blend(ts, *lap), ut, nu, dt)


def pde_residual_backward_timelevels(problem, ts, u, ux, lap, nu, dt, g_loss):
    """-> (d_u1, d_u0, d_ux1, d_ux0, d_lap1, d_lap0), float64 (d_ux*: None when ux is None)."""
    ut = (_f32(u[0]) - _f32(u[1])).astype(np.float32)
    us = time_blend(ts, *u)
    uxs = None if ux is None else time_blend(ts, *ux)
    d_u, d_ux, d_lap, d_ut = pde_residual_backward(problem, us, uxs, time_blend(ts, *lap), ut, nu, dt, g_loss)
    a, b = _blend_adjoint(ts, d_u, d_u.ndim)
    la, lb = _blend_adjoint(ts, d_lap, d_lap.ndim)
    xa, xb = (None, None) if d_ux is None else _blend_adjoint(ts, d_ux, d_ux.ndim)
    return a + d_ut, b - d_ut, xa, xb, la, lb


def ns_residual_loss_timelevels(ts, u, ux, wx, wxx, w, nu, dt):
    wt = (_f32(w[0]) - _f32(w[1])).astype(np.float32)
    return ns_residual_loss(time_blend(ts, *u), time_blend(ts, *ux), time_blend(ts, *wx), time_blend(ts, *wxx), wt, nu, dt)


def ns_residual_backward_timelevels(ts, u, ux, wx, wxx, w, nu, dt, g_loss):
    """-> gradients in the order (u1, u0, ux1, ux0, wx1, wx0, wxx1, wxx0, w1, w0), float64."""
    wt = (_f32(w[0]) - _f32(w[1])).astype(np.float32)
    gs = ns_residual_backward(time_blend(ts, *u), time_blend(ts, *ux), time_blend(ts, *wx), time_blend(ts, *wxx), wt,
                              nu, dt, g_loss)
    out = []
    for g in gs[:4]:
        out += list(_blend_adjoint(ts, g, g.ndim))
    return tuple(out) + (gs[4], -gs[4])


# ---- split / prune compaction ---------------------------------------------------
def compact_plan(P, keep=None, extend=None, m=1):
    """Parent row of every output row: kept rows in order, then m rows per extended row (model_pn.py:598-605 cat order)."""
    keep = np.ones(P, bool) if keep is None else np.asarray(keep).reshape(-1).astype(bool)
    ext = np.zeros(P, bool) if extend is None else np.asarray(extend).reshape(-1).astype(bool)
    kept = np.nonzero(keep)[0]
    return np.concatenate([kept, np.repeat(np.nonzero(ext)[0], m)]).astype(np.int64), int(kept.size), int(ext.sum())


def split(means, scaling, transforms, u, cov, indices):
    """Model.split, model_pn.py:578-610, with the flattened covariances [P, d(d+1)/2]: eigen-decomposition of the
    symmetric matrix (float64), displacement = |lambda_max| * v_max, children at mu - disp, mu + disp.  The sign of
    v_max is not defined by the reference (LAPACK's choice): compare the two children of a parent as a set
    (`children_sorted`)."""
    idx = np.asarray(indices).reshape(-1).astype(bool)
    P, D = means.shape
    src, n_keep, n_split = compact_plan(P, ~idx, idx, 2)
    out_means = means[src].astype(np.float64)
    par = src[n_keep:].reshape(-1, 2)[:, 0]
    if D == 1:
        disp = np.abs(cov[par].astype(np.float64)).reshape(-1, 1)
    else:
        c = cov[par].astype(np.float64)
        A = np.stack([np.stack([c[:, 0], c[:, 1]], -1), np.stack([c[:, 1], c[:, 2]], -1)], -2)
        w, v = np.linalg.eigh(A)
        k = np.argmax(np.abs(w), axis=-1)
        rows = np.arange(par.size)
        disp = np.abs(w[rows, k])[:, None] * v[rows, :, k]
    out_means[n_keep::2] -= disp
    out_means[n_keep + 1::2] += disp
    out_u = u[src].astype(np.float32)
    out_u[n_keep:] = out_u[n_keep:] / np.float32(2.0)
    return out_means.astype(np.float32), scaling[src], transforms[src], out_u, src, n_keep


def children_sorted(out_means, n_keep):
    """Rows >= n_keep grouped per parent with the two children in lexicographic order (sign-of-eigenvector free)."""
    ch = np.asarray(out_means, np.float64)[n_keep:].reshape(-1, 2, out_means.shape[-1])
    swap = (ch[:, 0, 0] > ch[:, 1, 0]) | ((ch[:, 0, 0] == ch[:, 1, 0]) & (ch[:, 0, -1] > ch[:, 1, -1]))
    ch[swap] = ch[swap][:, ::-1]
    return ch
