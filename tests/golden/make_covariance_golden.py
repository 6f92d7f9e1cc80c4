"""Generates tests/golden/covariance_golden.npz by IMPORTING the reference's gaussians.py (CPU, matplotlib stubbed)
and running build_covariances (gaussians.py:163-193) with torch.autograd for the gradients, on the parameter
recipes the drivers use (model_pn.py:460-461,495-497; test_derivatives.py:50-58).
Run:  python tests/golden/make_covariance_golden.py"""
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


def case(seed, P, base, spread, tspread):
    g = torch.Generator().manual_seed(seed)
    s = (base * torch.exp(spread * torch.randn(P, 2, generator=g))).requires_grad_(True)
    t = (tspread * torch.randn(P, 1, generator=g)).requires_grad_(True)
    cov, con = gaussians.build_covariances(s, t)
    g_cov = torch.randn(cov.shape, generator=g)
    g_con = torch.randn(con.shape, generator=g) / con.detach().abs().mean()  # comparable magnitudes of both terms
    ds, dt = torch.autograd.grad((cov * g_cov).sum() + (con * g_con).sum(), (s, t))
    return {k: v.detach().numpy().astype(np.float32) for k, v in
            dict(s=s, t=t, cov=cov, con=con, g_cov=g_cov, g_con=g_con, ds=ds, dt=dt).items()}


if __name__ == "__main__":
    out = {}
    cases = {"model_pn": (0, 400, np.exp(-4.0) * 20 / 1000.0, 0.5, 0.3),   # P = 1M recipe of SURVEY §8d
             "derivatives": (1, 300, np.exp(-3.5), 0.2, 0.5),                # test_derivatives.py-style
             "wide": (2, 200, 0.3, 1.0, 1.0)}
    for name, a in cases.items():
        for k, v in case(*a).items():
            out[name + "/" + k] = v
    np.savez_compressed(os.path.join(HERE, "covariance_golden.npz"), **out)
    print("wrote", len(out), "arrays")
