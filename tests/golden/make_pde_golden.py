"""Generates tests/golden/pde_golden.npz by IMPORTING the reference's model_pn.py in this container (CPU) and
calling its own `Model.pde_rhs` (model_pn.py:612-642) — unbound, on a stand-in `self` that carries only the two
attributes it reads (`problem`, `nu`) — followed by the loss lines of `compute_loss` (model_pn.py:822-841) and
torch.autograd for the gradients.  matplotlib is absent here and is stubbed; `diff_gaussian_sampling` resolves to
this repo's drop-in (imported, never called).  Run:  python tests/golden/make_pde_golden.py"""
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
import model_pn  # noqa: E402


def case(problem, seed, N, C, D, nu, dt):
    g = torch.Generator().manual_seed(seed)
    u = torch.randn(N, C, generator=g).requires_grad_(True)
    ux = torch.randn(N, D, C, generator=g).requires_grad_(True)
    uxx = torch.randn(N, D, D, C, generator=g).requires_grad_(True)
    ut = (0.1 * torch.randn(N, C, generator=g)).requires_grad_(True)
    me = types.SimpleNamespace(problem=problem, nu=nu)
    rhs = dt * model_pn.Model.pde_rhs(me, None, u, ux, uxx)            # model_pn.py:824
    if problem == model_pn.Problem.WAVE:                               # model_pn.py:835-837
        loss = 0.01 * torch.mean((ut[..., 0] - rhs[..., 0]) ** 2) + torch.mean((ut[..., 1] - rhs[..., 1]) ** 2)
    else:                                                              # model_pn.py:826-830
        loss = torch.mean((ut - rhs) ** 2)
    up = torch.randn((), generator=g) + 2.0
    gu, gux, guxx, gut = torch.autograd.grad(loss * up, (u, ux, uxx, ut), allow_unused=True)
    z = lambda t, like: torch.zeros_like(like) if t is None else t
    guxx = z(guxx, uxx)
    out = dict(problem=np.int32(problem.value), nu=np.float32(nu), dt=np.float32(dt), u=u, ux=ux, ut=ut,
               lap=uxx[:, 0, 0] + uxx[:, 1, 1] if D == 2 else uxx[:, 0, 0], residual=(ut - rhs), loss=loss, up=up,
               d_u=z(gu, u), d_ux=z(gux, ux), d_ut=z(gut, ut),
               # the loss depends on the Hessian only through its trace: d/dlap = the [0,0] entry of d/duxx
               d_lap=guxx[:, 0, 0])
    # ... and that entry equals the [1,1] one while the off-diagonal ones vanish (checked here, not stored)
    assert torch.equal(guxx[:, 0, 0], guxx[:, 1, 1]) and not guxx[:, 0, 1].any() and not guxx[:, 1, 0].any()
    return {k: (v.detach().numpy().astype(np.float32) if isinstance(v, torch.Tensor) else v) for k, v in out.items()}


def ns_case(seed, N, nu, dt):
    """NAVIER_STOKES the way Model.sample / compute_loss chain it (model_pn.py:772-781, 818, 824, 843-845): the
    sampler's C = 2 outputs of two time levels -> vorticity terms -> pde_rhs -> the two loss terms."""
    g = torch.Generator().manual_seed(seed)
    P = model_pn.Problem
    mk = lambda *sh: torch.randn(*sh, generator=g).requires_grad_(True)
    u, ux, uxx, uxxx = mk(N, 2), mk(N, 2, 2), mk(N, 2, 2, 2), mk(N, 2, 2, 2, 2)
    ux_prev = mk(N, 2, 2)
    w, w_prev = ux[:, 0, 1] - ux[:, 1, 0], ux_prev[:, 0, 1] - ux_prev[:, 1, 0]      # model_pn.py:779
    wx = uxx[..., 0, 1] - uxx[..., 1, 0]                                            # :780
    wxx = uxxx[..., 0, 1] - uxxx[..., 1, 0]                                         # :781
    for t in (wx, wxx):
        t.retain_grad()
    wt = (w - w_prev).detach().requires_grad_(True)                                 # :818 (a leaf here: d_wt is stored)
    me = types.SimpleNamespace(problem=P.NAVIER_STOKES, nu=nu)
    rhs = dt * model_pn.Model.pde_rhs(me, None, u, ux, uxx, wx, wxx)                # :824
    loss = torch.mean((ux[:, 0, 0] + ux[:, 1, 1]) ** 2) + torch.mean((wt - rhs) ** 2)   # :843-845
    up = torch.randn((), generator=g) + 2.0
    (loss * up).backward()
    out = dict(problem=np.int32(P.NAVIER_STOKES.value), nu=np.float32(nu), dt=np.float32(dt), u=u, ux=ux, uxx=uxx,
               uxxx=uxxx, ux_prev=ux_prev, w=w, w_prev=w_prev, wx=wx, wxx=wxx, wt=wt, residual=(wt - rhs),
               divergence=ux[:, 0, 0] + ux[:, 1, 1], loss=loss, up=up, d_u=u.grad, d_ux=ux.grad, d_wx=wx.grad,
               d_wxx=wxx.grad, d_wt=wt.grad, d_uxx=uxx.grad, d_uxxx=uxxx.grad)
    return {k: (v.detach().numpy().astype(np.float32) if isinstance(v, torch.Tensor) else v) for k, v in out.items()}


def main():
    P = model_pn.Problem
    cases = {
        "diffusion": case(P.DIFFUSION, 1, 1025, 1, 2, 0.0, 0.01),
        "diffusion_c3": case(P.DIFFUSION, 2, 1000, 3, 2, 0.0, 0.5),
        "burgers": case(P.BURGERS, 3, 1025, 1, 2, 0.01 / np.pi, 0.01),
        "burgers_c2": case(P.BURGERS, 4, 777, 2, 2, 0.3, 0.25),
        "wave": case(P.WAVE, 5, 1025, 2, 2, 0.0, 0.02),
        "navier_stokes": ns_case(6, 1025, 0.01, 0.02),
        "navier_stokes_b": ns_case(7, 333, 0.5, 0.25),
    }
    flat = {}
    for name, d in cases.items():
        for k, v in d.items():
            flat[name + "/" + k] = v
    np.savez_compressed(os.path.join(HERE, "pde_golden.npz"), **flat)
    print("wrote pde_golden.npz:", {n: float(d["loss"]) for n, d in cases.items()})


if __name__ == "__main__":
    main()
