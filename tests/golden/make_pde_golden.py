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


def _blend(rule, ts, cur, prev):
    """model_pn.py:794-813, the lines of compute_loss for the three integration rules."""
    if rule == "trapezoid":
        sh = (-1,) + (1,) * (cur.dim() - 1)
        return ts.reshape(sh) * cur + (1 - ts.reshape(sh)) * prev
    return prev if rule == "forward" else cur


def levels_case(problem, rule, seed, N, C, D, nu, dt):
    """compute_loss with two time levels (model_pn.py:794-837): blend -> ut = u[-1] - u[-2] -> dt * pde_rhs -> loss."""
    g = torch.Generator().manual_seed(seed)
    mk = lambda *sh: torch.randn(*sh, generator=g).requires_grad_(True)
    u1, u0, ux1, ux0, uxx1, uxx0 = mk(N, C), mk(N, C), mk(N, D, C), mk(N, D, C), mk(N, D, D, C), mk(N, D, D, C)
    ts = torch.rand(N, generator=g)
    u_s, ux, uxx = _blend(rule, ts, u1, u0), _blend(rule, ts, ux1, ux0), _blend(rule, ts, uxx1, uxx0)
    ut = u1 - u0                                                        # model_pn.py:820
    me = types.SimpleNamespace(problem=problem, nu=nu)
    rhs = dt * model_pn.Model.pde_rhs(me, None, u_s, ux, uxx)           # :824
    if problem == model_pn.Problem.WAVE:
        loss = 0.01 * torch.mean((ut[..., 0] - rhs[..., 0]) ** 2) + torch.mean((ut[..., 1] - rhs[..., 1]) ** 2)
    else:
        loss = torch.mean((ut - rhs) ** 2)
    up = torch.randn((), generator=g) + 2.0
    leaves = (u1, u0, ux1, ux0, uxx1, uxx0)
    grads = [torch.zeros_like(t) if gr is None else gr
             for gr, t in zip(torch.autograd.grad(loss * up, leaves, allow_unused=True), leaves)]
    tr = lambda h: h[:, 0, 0] + h[:, 1, 1]
    for gh in grads[4:]:  # the loss sees the Hessians only through their traces
        assert torch.equal(gh[:, 0, 0], gh[:, 1, 1]) and not gh[:, 0, 1].any() and not gh[:, 1, 0].any()
    out = dict(problem=np.int32(problem.value), rule=np.int32({"trapezoid": 2, "forward": 0, "backward": 1}[rule]),
               nu=np.float32(nu), dt=np.float32(dt), ts=ts, u1=u1, u0=u0, ux1=ux1, ux0=ux0, lap1=tr(uxx1), lap0=tr(uxx0),
               residual=ut - rhs, loss=loss, up=up, d_u1=grads[0], d_u0=grads[1], d_ux1=grads[2], d_ux0=grads[3],
               d_lap1=grads[4][:, 0, 0], d_lap0=grads[5][:, 0, 0])
    return {k: (v.detach().numpy().astype(np.float32) if isinstance(v, torch.Tensor) else v) for k, v in out.items()}


def ns_levels_case(seed, N, nu, dt):
    """NAVIER_STOKES, TRAPEZOID rule: model_pn.py:794-805 blends, :818 wt, :824 rhs, :843-845 loss."""
    g = torch.Generator().manual_seed(seed)
    P = model_pn.Problem
    mk = lambda *sh: torch.randn(*sh, generator=g).requires_grad_(True)
    names = ("u1", "u0", "ux1", "ux0", "wx1", "wx0", "wxx1", "wxx0", "w1", "w0")
    shapes = ((N, 2),) * 2 + ((N, 2, 2),) * 2 + ((N, 2),) * 2 + ((N, 2, 2),) * 2 + ((N,),) * 2
    t = {k: mk(*sh) for k, sh in zip(names, shapes)}
    ts = torch.rand(N, generator=g)
    u = _blend("trapezoid", ts, t["u1"], t["u0"])
    ux = _blend("trapezoid", ts, t["ux1"], t["ux0"])
    wx = _blend("trapezoid", ts, t["wx1"], t["wx0"])
    wxx = _blend("trapezoid", ts, t["wxx1"], t["wxx0"])
    wt = t["w1"] - t["w0"]
    me = types.SimpleNamespace(problem=P.NAVIER_STOKES, nu=nu)
    rhs = dt * model_pn.Model.pde_rhs(me, None, u, ux, None, wx, wxx)
    loss = torch.mean((ux[:, 0, 0] + ux[:, 1, 1]) ** 2) + torch.mean((wt - rhs) ** 2)
    up = torch.randn((), generator=g) + 2.0
    grads = torch.autograd.grad(loss * up, [t[k] for k in names])
    out = dict(nu=np.float32(nu), dt=np.float32(dt), ts=ts, residual=wt - rhs, divergence=ux[:, 0, 0] + ux[:, 1, 1],
               loss=loss, up=up, **t, **{"d_" + k: gr for k, gr in zip(names, grads)})
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
        "levels_diffusion": levels_case(P.DIFFUSION, "trapezoid", 11, 1025, 1, 2, 0.0, 0.01),
        "levels_burgers": levels_case(P.BURGERS, "trapezoid", 12, 1025, 1, 2, 0.01 / np.pi, 0.01),
        "levels_burgers_c2": levels_case(P.BURGERS, "trapezoid", 13, 777, 2, 2, 0.3, 0.25),
        "levels_burgers_forward": levels_case(P.BURGERS, "forward", 14, 500, 1, 2, 0.2, 0.1),
        "levels_burgers_backward": levels_case(P.BURGERS, "backward", 15, 500, 1, 2, 0.2, 0.1),
        "levels_wave": levels_case(P.WAVE, "trapezoid", 16, 1025, 2, 2, 0.0, 0.02),
        "levels_navier_stokes": ns_levels_case(17, 1025, 0.05, 0.05),
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
