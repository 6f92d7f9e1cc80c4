"""Generates tests/golden/densify_golden.npz by IMPORTING the reference's model_pn.py in this container (CPU) and
calling its own `Model.split` (model_pn.py:578-610) — unbound, on a stand-in `self` carrying the attributes it reads
and writes.  torch.zeros / torch.ones are wrapped to ignore the hard-wired device="cuda" (model_pn.py:603-605);
matplotlib is stubbed.  Run:  python tests/golden/make_densify_golden.py"""
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
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, "/root/reference")
import gaussians  # noqa: E402
import model_pn  # noqa: E402

_zeros, _ones = torch.zeros, torch.ones


def _nodev(f):
    def g(*a, **k):
        k.pop("device", None)
        return f(*a, **k)
    return g


def case(seed, P, C, frac):
    g = torch.Generator().manual_seed(seed)
    means = torch.rand(P, 2, generator=g) * 2 - 1
    scaling = torch.exp(torch.randn(P, 2, generator=g) * 0.5 - 4.0)
    scaling[:, 1] *= 3.0  # clearly anisotropic: the major axis is well defined
    transforms = torch.randn(P, 1, generator=g) * 0.6
    u = torch.rand(P, C, generator=g) * 2 - 1
    indices = torch.rand(P, generator=g) < frac
    full_cov, full_con = gaussians.build_full_covariances(scaling, transforms)
    cov, _ = gaussians.flatten_covariances(full_cov, full_con)
    me = types.SimpleNamespace(d=2, split_size=2, means=means.clone(), scaling=scaling.clone(),
                               transforms=transforms.clone(), u=u.clone(), boundaries=torch.zeros(P, 1),
                               boundary_mask=torch.ones(P, 1, dtype=torch.bool), full_covariances=full_cov)
    torch.zeros, torch.ones = _nodev(_zeros), _nodev(_ones)
    try:
        model_pn.Model.split(me, indices)
    finally:
        torch.zeros, torch.ones = _zeros, _ones
    out = dict(means=means, scaling=scaling, transforms=transforms, u=u, cov=cov, indices=indices.to(torch.uint8),
               out_means=me.means, out_scaling=me.scaling, out_transforms=me.transforms, out_u=me.u)
    return {k: v.detach().numpy() for k, v in out.items()}


def main():
    cases = {"small": case(1, 257, 1, 0.2), "c2": case(2, 1500, 2, 0.05), "all": case(3, 64, 1, 1.1), "none": case(4, 64, 1, -1.0)}
    flat = {}
    for name, d in cases.items():
        for k, v in d.items():
            flat[name + "/" + k] = v
    np.savez_compressed(os.path.join(HERE, "densify_golden.npz"), **flat)
    print({n: (d["means"].shape[0], d["out_means"].shape[0]) for n, d in cases.items()})


if __name__ == "__main__":
    main()
