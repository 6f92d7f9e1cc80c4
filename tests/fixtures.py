"""Seeded input recipes shared by the parity tests, the golden-vector generators and bench.py.

Shapes follow the reference's own driver scripts (cited per recipe; /root/reference/...), re-stated in
numpy so the same inputs exist on the CPU (oracle) and on the GPU.  Every recipe returns a dict with
float32 arrays: means [P,D], values [P,C], covariances [P,K], conics [P,K], samples [N,D]
(K = D(D+1)/2, flattened [xx, xy, yy] as gaussians.py:186-189 does).
"""
import numpy as np


def covariances_from(s, t):
    """gaussians.py:163-193 build_covariances: Sigma = [[s0, tau],[tau, s1]], tau = tanh(t) * sqrt(s0 s1);
    conic = Sigma^-1; both flattened [xx, xy, yy].  (float64 maths, rounded to float32 at the end.)"""
    s = np.asarray(s, np.float64)
    d = s.shape[-1]
    if d == 1:
        return s.astype(np.float32), (1.0 / s).astype(np.float32)
    tau = np.tanh(np.asarray(t, np.float64).reshape(-1)) * np.sqrt(s[:, 0] * s[:, 1])
    det = s[:, 0] * s[:, 1] - tau * tau
    cov = np.stack([s[:, 0], tau, s[:, 1]], 1)
    con = np.stack([s[:, 1] / det, -tau / det, s[:, 0] / det], 1)
    return cov.astype(np.float32), con.astype(np.float32)


def _grid_samples(res, scale=1.0):
    t = np.linspace(-1, 1, res, dtype=np.float32) * np.float32(scale)
    gx, gy = np.meshgrid(t, t, indexing="xy")
    return np.stack([gx, gy], -1).reshape(res * res, 2).astype(np.float32)


def _pack(means, values, cov, con, samples):
    return dict(means=np.ascontiguousarray(means, np.float32), values=np.ascontiguousarray(values, np.float32),
                covariances=np.ascontiguousarray(cov, np.float32), conics=np.ascontiguousarray(con, np.float32),
                samples=np.ascontiguousarray(samples, np.float32))


def grid10(res=128):
    """test_gaussian_sampling.py:13-46 — 10x10 grid of isotropic sigma^2 = e^-4 Gaussians, res^2 image samples."""
    t = np.linspace(-1, 1, 10, dtype=np.float32)
    gx, gy = np.meshgrid(t, t, indexing="ij")
    means = np.stack([gx, gy], -1).reshape(100, 2)
    d = means.astype(np.float64) - np.array([0.1, 0.4])
    values = np.exp(-0.5 * 10.0 * (d * d).sum(-1)).reshape(100, 1)
    cov, con = covariances_from(np.full((100, 2), np.exp(-4.0)), np.zeros(100))
    return _pack(means, values, cov, con, _grid_samples(res))


def density(seed=0, n=12, res=96):
    """test_density.py:11-35 — jittered grid, v = 0.5."""
    rng = np.random.default_rng(seed)
    t = np.linspace(-1, 1, n)
    gx, gy = np.meshgrid(t, t, indexing="ij")
    means = np.stack([gx, gy], -1).reshape(n * n, 2) + rng.uniform(-1, 1, (n * n, 2)) * (1.0 / n)
    means = np.clip(means, -1, 1)
    cov, con = covariances_from(np.exp(rng.normal(-4.0, 0.3, (n * n, 2))), np.zeros(n * n))
    return _pack(means, np.full((n * n, 1), 0.5), cov, con, _grid_samples(res))


def torus(res=96):
    """test_torus.py:11-30 — a column of Gaussians at x = -0.95, sigma^2 = e^-3: wraps across the x = +-1 seam."""
    n = 10
    means = np.stack([np.full(n, -0.95), np.linspace(-1, 1, n)], -1)
    cov, con = covariances_from(np.full((n, 2), np.exp(-3.0)), np.zeros(n))
    return _pack(means, np.full((n, 1), 0.5), cov, con, _grid_samples(res))


def line1d(n=20, res=200):
    """test_1d.py:11-27 — 1-D, sigma^2 = e^-5, samples given as a 1-D array."""
    means = np.linspace(-1, 1, n).reshape(n, 1)
    values = np.exp(-(means * 4) ** 2).reshape(n, 1)
    cov = np.full((n, 1), np.exp(-5.0))
    return _pack(means, values, cov, 1.0 / cov, np.linspace(-1, 1, res).reshape(res, 1))


def derivatives(seed=0, res=64):
    """test_derivatives.py:13-70 — off-diagonal 0.5*tanh, sigma^2 = e^-3.5, 64^2 points, random values."""
    rng = np.random.default_rng(seed)
    n = 5
    t = np.linspace(-1, 1, n)
    gx, gy = np.meshgrid(t, t, indexing="ij")
    means = np.stack([gx, gy], -1).reshape(n * n, 2) * 0.8
    cov, con = covariances_from(np.full((n * n, 2), np.exp(-3.5)) * np.exp(rng.normal(0, 0.2, (n * n, 2))),
                                rng.normal(0, 0.5, n * n))
    return _pack(means, rng.uniform(-1, 1, (n * n, 1)), cov, con, _grid_samples(res))


def synthetic(P, N, seed=0, C=1, D=2, edge_fraction=0.0, anisotropy=1.0, scale=1.0):
    """SURVEY.md §8d recipe (after model_pn.py:439-502 randomize): jittered ceil(sqrt(P))^2 grid truncated to P,
    s = e^-4 * 20/sqrt(P) * exp(0.5 N(0,1)) per axis, off-diagonal t = 0.3 N(0,1), values U(-1,1), points
    U(-1,1)^D plus the domain corners so the tile grid is exactly 4^D.  edge_fraction moves that share of the
    means to within 3 sigma of the +-1 edges (config C4); anisotropy multiplies s1."""
    rng = np.random.default_rng(seed)
    if D == 2:
        n = int(np.ceil(np.sqrt(P)))
        t = np.linspace(-1, 1, n)
        gx, gy = np.meshgrid(t, t, indexing="ij")
        means = np.stack([gx, gy], -1).reshape(n * n, 2)
        means = means + rng.uniform(-1, 1, means.shape) * 0.5 * (2.0 / max(n - 1, 1))
        means = np.clip(means[:P], -1, 1)
        s = np.exp(-4.0) * 20.0 / np.sqrt(P) * np.exp(0.5 * rng.normal(size=(P, 2)))
        s[:, 1] *= anisotropy
        tt = 0.3 * rng.normal(size=P)
        if edge_fraction > 0:
            k = int(P * edge_fraction)
            idx = rng.choice(P, k, replace=False)
            side = rng.integers(0, 2, k) * 2 - 1
            means[idx, 0] = side * (1.0 - rng.uniform(0, 3, k) * np.sqrt(s[idx, 0]))
            means = np.clip(means, -1, 1)
        cov, con = covariances_from(s, tt)
        samples = rng.uniform(-1, 1, (N - 4, 2))
        samples = np.concatenate([samples, np.array([[-1, -1], [1, 1], [-1, 1], [1, -1]], np.float64)], 0)
    else:
        means = np.linspace(-1, 1, P).reshape(P, 1) + rng.uniform(-1, 1, (P, 1)) * (0.5 * 2.0 / max(P - 1, 1))
        means = np.clip(means, -1, 1)
        s = np.exp(-4.0) * 20.0 / P * np.exp(0.5 * rng.normal(size=(P, 1)))
        cov, con = covariances_from(s, None)
        samples = np.concatenate([rng.uniform(-1, 1, (N - 2, 1)), np.array([[-1.0], [1.0]])], 0)
    values = rng.uniform(-1, 1, (P, C))
    return _pack(means * scale, values, cov * scale * scale, con / (scale * scale), samples * scale)


def wide(seed=0, P=40, N=3000):
    """A handful of LARGE Gaussians (r > tile width): mixed torus wrap inside tiles, rect clamps to the whole grid."""
    rng = np.random.default_rng(seed)
    means = rng.uniform(-1, 1, (P, 2))
    cov, con = covariances_from(np.exp(rng.uniform(-4.5, -1.0, (P, 2))), rng.normal(0, 0.4, P))
    samples = np.concatenate([rng.uniform(-1, 1, (N - 2, 2)), np.array([[-1.0, -1.0], [1.0, 1.0]])], 0)
    return _pack(means, rng.uniform(-1, 1, (P, 2)), cov, con, samples)


RECIPES = {
    "grid10": lambda: grid10(96),
    "density": density,
    "torus": torus,
    "line1d": line1d,
    "derivatives": derivatives,
    "synth2d": lambda: synthetic(3000, 20000, seed=1),
    "synth2d_c2": lambda: synthetic(1500, 8000, seed=2, C=2),
    "synth2d_c3": lambda: synthetic(600, 4000, seed=5, C=3),
    "synth1d": lambda: synthetic(500, 5000, seed=3, D=1),
    "synth1d_c2": lambda: synthetic(300, 3000, seed=4, D=1, C=2),
    "edge_aniso": lambda: synthetic(2000, 12000, seed=6, edge_fraction=0.1, anisotropy=4.0),
    "wide": wide,
    "scaled": lambda: synthetic(800, 6000, seed=7, scale=2.5),
}


# ---- neighbour aggregation (aggregate_neighbors.cu) -------------------------------------------------
def agg_params(P, L, K, E, D, seed=0, dt_rows=None):
    """test_neighbor_aggregation.py:69-80 / model_pn.py:199-230: features, transform [L,L], queries/keys [P,K],
    frequencies [(E-1)//D//2] = 10 N(0,1), distance_transform [dt_rows or L, 2E] (only row 0 is read by the kernel,
    aggregate_neighbors.cu:185-193)."""
    rng = np.random.default_rng(seed)
    f = lambda *s: rng.uniform(-1, 1, s).astype(np.float32)
    return dict(features=f(P, L), transform=f(L, L), queries=f(P, K), keys=f(P, K),
                frequencies=(rng.normal(size=(E - 1) // D // 2) * 10).astype(np.float32),
                distance_transform=f(dt_rows or L, 2 * E))


def agg_grid5():
    """test_neighbor_aggregation.py:15-27,69-80 (gradcheck branch): 5x5 grid, sigma^2 = e^-1.5, L=2, K=4, E=21."""
    t = np.linspace(-1, 1, 5)
    gx, gy = np.meshgrid(t, t, indexing="ij")
    means = np.stack([gx, gy], -1).reshape(25, 2)
    cov, con = covariances_from(np.full((25, 2), np.exp(-1.5)), np.zeros(25))
    values = np.zeros((25, 1))
    values[:5] = 1.0
    rec = _pack(means, values, cov, con, _grid_samples(64))
    rec.update(agg_params(25, 2, 4, 21, 2, seed=11))
    return rec


def agg_model(P=700, seed=12, D=2, **kw):
    """model_pn.py call pattern (samples = means, model_pn.py:648): L = K = 16, E = 25, nF = 6 (1-D: 12)."""
    rec = synthetic(P, 64, seed=seed, D=D, **kw)
    rec["samples"] = rec["means"].copy()
    rec.update(agg_params(P, 16, 16, 25, D, seed=seed + 100))
    return rec


AGG_RECIPES = {
    "agg_grid5": agg_grid5,
    "agg_model": agg_model,
    "agg_model_edge": lambda: agg_model(900, seed=13, edge_fraction=0.15, anisotropy=3.0),
    "agg_model_1d": lambda: agg_model(300, seed=14, D=1),
    "agg_scaled": lambda: agg_model(400, seed=15, scale=2.5),
}
