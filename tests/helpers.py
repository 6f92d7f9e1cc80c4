"""Shared helpers for the parity tests (GPU side goes through pigs_b200's public API / C ABI)."""
import numpy as np

from oracle import pigs_oracle as po

TOL = 1e-5  # BASELINE.json north_star: 1e-5 relative (norm-wise: max|a-b| <= TOL * max|b|)


def rel_err(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    den = np.abs(b).max()
    if den == 0:
        return float(np.abs(a).max())
    return float(np.abs(a - b).max() / den)


def to_dev(rec, device="cuda"):
    import torch
    return {k: torch.from_numpy(v).to(device) for k, v in rec.items()}


def oracle_state(rec, recip_mode=1):
    return po.preprocess(rec["means"], rec["covariances"], rec["samples"], recip_mode=recip_mode)


def upstream(rec, fn, seed=0):
    """seeded N(0,1) upstream gradient for function index fn (0 value .. 3 third)"""
    N, D = rec["samples"].shape
    C = rec["values"].shape[1]
    rng = np.random.default_rng(1000 + seed + fn)
    return rng.normal(size=po.out_shape(fn, N, D, C)).astype(np.float32)


def ranges_np(t, blocks):
    """uint8 tensor (blocks*8+8 bytes) -> [(blocks+1), 2] uint32"""
    return t.cpu().numpy().view(np.uint32).reshape(blocks + 1, 2)


def agg_upstream(rec, seed=0):
    """seeded N(0,1) upstream gradient of aggregate_neighbors' output [P, L]"""
    rng = np.random.default_rng(2000 + seed)
    return rng.normal(size=rec["features"].shape).astype(np.float32)


AGG_KEYS = ("features", "transform", "queries", "keys", "frequencies", "distance_transform")
# shared-parameter gradients vs the reference CUDA golden: bounded by the reference's own fp32-atomic noise
AGG_SHARED_TOL = 1e-4
