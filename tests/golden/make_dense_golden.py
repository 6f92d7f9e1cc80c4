"""Generates tests/golden/dense_golden.npz by IMPORTING the reference's pure-PyTorch path
(/root/reference/gaussians.py:48-58 sample_gaussians, :89-101 gaussian_derivative, :103-116 gaussian_derivative2)
in this container (CPU).  matplotlib is absent here and is stubbed in sys.modules; gaussian_derivative2 builds a
dead `torch.ones(..., device="cuda")` (gaussians.py:110), so torch.ones is wrapped to ignore the device argument
for the duration of that call.  Run:  python tests/golden/make_dense_golden.py
Fixtures: Gaussians wide enough that every tile rect covers the whole (tiny) sample domain, so the reference CUDA
pair set equals the dense one and no pair wraps — the only regime where gaussians.py is an exact oracle."""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
for m in ("matplotlib", "matplotlib.pyplot", "matplotlib.transforms", "matplotlib.patches"):
    sys.modules[m] = types.ModuleType(m)
sys.modules["matplotlib.transforms"].Affine2D = object
sys.modules["matplotlib.patches"].Ellipse = object
sys.path.insert(0, "/root/reference")
import gaussians  # noqa: E402

_ones = torch.ones


def _ones_cpu(*a, **k):
    k.pop("device", None)
    return _ones(*a, **k)


def case(seed, P, N, C, span):
    g = torch.Generator().manual_seed(seed)
    means = (torch.rand(P, 2, generator=g) - 0.5) * span
    s = torch.exp(torch.randn(P, 2, generator=g) * 0.2 - 2.0)
    t = torch.randn(P, 1, generator=g) * 0.4
    full_cov, full_con = gaussians.build_full_covariances(s, t)
    cov, con = gaussians.flatten_covariances(full_cov, full_con)
    values = torch.rand(P, C, generator=g) * 2 - 1
    samples = (torch.rand(N, 2, generator=g) - 0.5) * span
    out = dict(means=means, conics=con, cov=cov, values=values, samples=samples)
    m = means.clone().requires_grad_(True)
    fc = full_con.clone().requires_grad_(True)
    v = values.clone().requires_grad_(True)
    torch.ones = _ones_cpu
    try:
        fns = dict(u=gaussians.sample_gaussians, du=gaussians.gaussian_derivative, ddu=gaussians.gaussian_derivative2)
        for key, fn in fns.items():
            o = fn(m, fc, v, samples)
            up = torch.randn(o.shape, generator=g)
            gm, gc, gv = torch.autograd.grad((o * up).sum(), (m, fc, v))
            out[key] = o.detach()
            out["g" + key] = up
            out[key + "_dmeans"] = gm
            out[key + "_dvalues"] = gv
            # flattened [xx, xy, yy] parameterisation: the stored xy entry feeds both off-diagonals
            out[key + "_dconics"] = torch.stack([gc[:, 0, 0], gc[:, 0, 1] + gc[:, 1, 0], gc[:, 1, 1]], 1)
    finally:
        torch.ones = _ones
    return {k: x.detach().numpy().astype(np.float32) for k, x in out.items()}


def case_full(seed, P, N, C, D, asym=0.0):
    """Dense path proper: FULL [n, d, d] conics as the functions take them, gradients w.r.t. the full
    matrices, 1-D and 2-D, samples anywhere (no tile / torus regime restriction).  asym > 0 perturbs the conics
    off-diagonals independently: the functions do not assume symmetry."""
    g = torch.Generator().manual_seed(seed)
    means = torch.rand(P, D, generator=g) * 2 - 1
    s = torch.exp(torch.randn(P, D, generator=g) * 0.3 - 3.0)
    t = torch.randn(P, 1, generator=g) * 0.4
    if D == 2:
        _, full_con = gaussians.build_full_covariances(s, t)
        if asym:
            full_con = full_con + asym * full_con.abs().mean() * torch.randn(P, 2, 2, generator=g)
    else:
        full_con = (1.0 / s).reshape(P, 1, 1)
    values = torch.rand(P, C, generator=g) * 2 - 1
    samples = torch.rand(N, D, generator=g) * 2 - 1
    out = dict(means=means, conics=full_con, values=values, samples=samples)
    m, fc, v = (x.clone().requires_grad_(True) for x in (means, full_con, values))
    torch.ones = _ones_cpu
    try:
        fns = dict(u=gaussians.sample_gaussians, du=gaussians.gaussian_derivative, ddu=gaussians.gaussian_derivative2)
        for key, fn in fns.items():
            o = fn(m, fc, v, samples)
            up = torch.randn(o.shape, generator=g)
            gm, gc, gv = torch.autograd.grad((o * up).sum(), (m, fc, v))
            out.update({key: o.detach(), "g" + key: up, key + "_dmeans": gm, key + "_dconics": gc, key + "_dvalues": gv})
    finally:
        torch.ones = _ones
    return {k: x.detach().numpy().astype(np.float32) for k, x in out.items()}


if __name__ == "__main__":
    full = {}
    for tag, args in (("f2", (10, 300, 2500, 1, 2)), ("f2c3", (11, 130, 700, 3, 2)), ("f2asym", (12, 200, 900, 2, 2, 0.05)),
                      ("f1", (13, 257, 3000, 1, 1)), ("f1c2", (14, 64, 500, 2, 1))):
        for k, x in case_full(*args).items():
            full[f"{tag}_{k}"] = x
    np.savez_compressed(os.path.join(HERE, "dense_full_golden.npz"), **full)
    print("wrote dense_full_golden.npz", {k: v.shape for k, v in full.items() if k.startswith("f2_")})
    res = {}
    for tag, args in (("a", (0, 40, 300, 1, 0.4)), ("b", (1, 25, 200, 2, 0.3))):
        for k, x in case(*args).items():
            res[f"{tag}_{k}"] = x
    np.savez_compressed(os.path.join(HERE, "dense_golden.npz"), **res)
    print("wrote", os.path.join(HERE, "dense_golden.npz"), {k: v.shape for k, v in res.items() if k.startswith("a_")})
