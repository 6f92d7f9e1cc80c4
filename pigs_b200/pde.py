"""PDE right-hand side + residual loss of kr4b/pigs' `model_pn.py` as ONE fused CUDA pass each way (SURVEY.md §8f
rank 2), fed by the sampler's Laplacian output (`GaussianSampler.sample_laplacian()`, the trace `uxx[:,0,0] +
uxx[:,1,1]` the reference takes of its [N, D, D, C] Hessian):

    model_pn.py:612-625  pde_rhs:   DIFFUSION lap | BURGERS nu*lap - u*ux[:,0] | WAVE (u1, 10*lap0 - 0.1*u1)
    model_pn.py:822-834  rhs = dt * pde_rhs(...);  pde_loss += mean((ut - rhs)**2)   (WAVE: 0.01*mean + mean)

    loss = pde_residual_loss(Problem.BURGERS, u, ux, lap, ut, dt, nu)     # 0-dim tensor, differentiable

    model_pn.py:629-631  pde_rhs:   NAVIER_STOKES nu*(wxx00 + wxx11) - (u0*wx0 + u1*wx1)   (vorticity form)
    model_pn.py:843-845  pde_loss += mean((ux[:,0,0] + ux[:,1,1])**2);  pde_loss += mean((wt - rhs)**2)
    model_pn.py:779-781  w = ux[:,0,1] - ux[:,1,0];  wx = uxx[...,0,1] - uxx[...,1,0];  wxx = uxxx[...,0,1] - uxxx[...,1,0]

    loss = navier_stokes_residual_loss(u, ux, wx, wxx, wt, dt, nu)       # 0-dim tensor, differentiable
    w, wx, wxx = vorticity(ux), vorticity(uxx), vorticity(uxxx)          # differentiable

    model_pn.py:794-811  TRAPEZOID x = ts*x[-1] + (1-ts)*x[-2] | FORWARD x[-2] | BACKWARD x[-1];  :820 ut = u[-1] - u[-2]

    loss = pde_residual_loss_timelevels(problem, time_samples, (u[-1], u[-2]), (ux[-1], ux[-2]), (lap[-1], lap[-2]), dt, nu)
    loss = navier_stokes_residual_loss_timelevels(time_samples, (u..), (ux..), (wx..), (wxx..), (w[-1], w[-2]), dt, nu)
    # time_samples: [N] tensor (TRAPEZOID) or 0.0 (FORWARD rule) / 1.0 (BACKWARD rule); the blends happen inside the kernel

A maintainer switches `compute_loss` to it by replacing the two lines above.  POISSON (its
reference code reads an undefined `t`, model_pn.py:620) is left on torch.
No CPU path: CUDA fp32 tensors only.
"""
import enum

import torch

from . import _C


class Problem(enum.Enum):
    """model_pn.py:13-19 (same member order, so the values agree)."""
    DIFFUSION = enum.auto()
    POISSON = enum.auto()
    BURGERS = enum.auto()
    WAVE = enum.auto()
    NAVIER_STOKES = enum.auto()
    TEST = enum.auto()


_NATIVE = {Problem.DIFFUSION.value: _C.PDE_DIFFUSION, Problem.BURGERS.value: _C.PDE_BURGERS, Problem.WAVE.value: _C.PDE_WAVE}


def _code(problem):
    v = problem.value if isinstance(problem, enum.Enum) else int(problem)
    if v not in _NATIVE:
        raise ValueError("Unexpected PDE problem for the fused residual: %r" % (problem,))
    return _NATIVE[v]


class _ResidualLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, code, u, ux, lap, ut, nu, dt):
        loss, _ = _C.pde_residual_forward(code, u, ux, lap, ut, nu, dt)
        ctx.code, ctx.nu, ctx.dt = code, nu, dt
        ctx.save_for_backward(*[t for t in (u, ux, lap, ut) if t is not None])
        ctx.has = (u is not None, ux is not None)
        return loss

    @staticmethod
    def backward(ctx, g):
        it = iter(ctx.saved_tensors)
        u = next(it) if ctx.has[0] else None
        ux = next(it) if ctx.has[1] else None
        lap, ut = next(it), next(it)
        need = (ctx.needs_input_grad[1], ctx.needs_input_grad[2], ctx.needs_input_grad[3], ctx.needs_input_grad[4])
        d_u, d_ux, d_lap, d_ut = _C.pde_residual_backward(ctx.code, u, ux, lap, ut, ctx.nu, ctx.dt, g, need)

        def shaped(d, like):
            return None if d is None else d.reshape(like.shape)

        return (None, shaped(d_u, u) if u is not None else None, shaped(d_ux, ux) if ux is not None else None,
                shaped(d_lap, lap), shaped(d_ut, ut), None, None)


def pde_residual_loss(problem, u, ux, lap, ut, dt, nu=0.0):
    """mean((ut - dt * pde_rhs)**2) for DIFFUSION / BURGERS, 0.01*mean(r0**2) + mean(r1**2) for WAVE
    (model_pn.py:822-841).  u, lap, ut: [N, C]; ux: [N, D, C] (may be None for DIFFUSION / WAVE)."""
    code = _code(problem)
    if code != _C.PDE_BURGERS:
        ux = None
    if code == _C.PDE_DIFFUSION:
        u = None
    return _ResidualLoss.apply(code, u, ux, lap, ut, float(nu), float(dt))


def pde_residual(problem, u, ux, lap, ut, dt, nu=0.0):
    """(no autograd) the residual ut - dt * pde_rhs itself, [N, C], and its loss."""
    code = _code(problem)
    loss, res = _C.pde_residual_forward(code, None if code == _C.PDE_DIFFUSION else u,
                                        ux if code == _C.PDE_BURGERS else None, lap, ut, nu, dt, want_residual=True)
    return res, loss


class _NsResidualLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, u, ux, wx, wxx, wt, nu, dt):
        loss, _, _ = _C.pde_ns_residual_forward(u, ux, wx, wxx, wt, nu, dt)
        ctx.nu, ctx.dt = nu, dt
        ctx.save_for_backward(u, ux, wx, wxx, wt)
        return loss

    @staticmethod
    def backward(ctx, g):
        saved = ctx.saved_tensors
        grads = _C.pde_ns_residual_backward(*saved, ctx.nu, ctx.dt, g, tuple(ctx.needs_input_grad[:5]))
        return tuple(None if d is None else d.reshape(t.shape) for d, t in zip(grads, saved)) + (None, None)


def navier_stokes_residual_loss(u, ux, wx, wxx, wt, dt, nu):
    """mean((ux[:,0,0] + ux[:,1,1])**2) + mean((wt - dt * (nu*(wxx00 + wxx11) - (u0*wx0 + u1*wx1)))**2)
    (model_pn.py:629-631, 824, 843-845).  u [N,2], ux [N,2,2], wx [N,2], wxx [N,2,2], wt [N]."""
    return _NsResidualLoss.apply(u, ux, wx, wxx, wt, float(nu), float(dt))


def navier_stokes_residual(u, ux, wx, wxx, wt, dt, nu):
    """(no autograd) -> (residual wt - dt * rhs [N], divergence ux00 + ux11 [N], loss)."""
    loss, res, div = _C.pde_ns_residual_forward(u, ux, wx, wxx, wt, nu, dt, want_residual=True)
    return res, div, loss


class _Vorticity(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        return _C.vorticity_forward(x)

    @staticmethod
    def backward(ctx, g):
        return _C.vorticity_backward(g)


def vorticity(x):
    """x[..., 0, 1] - x[..., 1, 0] over the trailing (spatial axis, channel) axes of a C = 2 sampler output
    (model_pn.py:779-781: ux -> w [N], uxx -> wx [N,2], uxxx -> wxx [N,2,2])."""
    return _Vorticity.apply(x)


class _ResidualLossLevels(torch.autograd.Function):
    @staticmethod
    def forward(ctx, code, ts, nu, dt, u1, u0, ux1, ux0, lap1, lap0):
        ux = None if ux1 is None else (ux1, ux0)
        loss, _ = _C.pde_residual_timelevels_forward(code, ts, (u1, u0), ux, (lap1, lap0), nu, dt)
        ctx.code, ctx.ts, ctx.nu, ctx.dt, ctx.has_ux = code, ts, nu, dt, ux is not None
        ctx.save_for_backward(*[t for t in (u1, u0, ux1, ux0, lap1, lap0) if t is not None])
        return loss

    @staticmethod
    def backward(ctx, g):
        it = iter(ctx.saved_tensors)
        u1, u0 = next(it), next(it)
        ux1, ux0 = (next(it), next(it)) if ctx.has_ux else (None, None)
        lap1, lap0 = next(it), next(it)
        ux = (ux1, ux0) if ctx.has_ux else None
        grads = _C.pde_residual_timelevels_backward(ctx.code, ctx.ts, (u1, u0), ux, (lap1, lap0), ctx.nu, ctx.dt, g,
                                                    tuple(ctx.needs_input_grad[4:10]))
        like = (u1, u0, ux1, ux0, lap1, lap0)
        return (None,) * 4 + tuple(None if d is None else d.reshape(t.shape) for d, t in zip(grads, like))


def pde_residual_loss_timelevels(problem, time_samples, u, ux, lap, dt, nu=0.0):
    """compute_loss's chain with the time integration inside the kernel (model_pn.py:794-837): u, ux, lap are pairs
    (level [-1], level [-2]) of the sampler's outputs (lap = `sample_laplacian()`), time_samples a [N] tensor
    (TRAPEZOID) or 0.0 / 1.0 (FORWARD / BACKWARD rule).  ux may be None unless BURGERS.  time_samples gets no gradient."""
    code = _code(problem)
    ux1, ux0 = ux if (ux is not None and code == _C.PDE_BURGERS) else (None, None)
    return _ResidualLossLevels.apply(code, time_samples, float(nu), float(dt), u[0], u[1], ux1, ux0, lap[0], lap[1])


class _NsResidualLossLevels(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ts, nu, dt, *levels):  # levels: u1, u0, ux1, ux0, wx1, wx0, wxx1, wxx0, w1, w0
        pairs = [levels[i:i + 2] for i in range(0, 10, 2)]
        loss, _, _ = _C.pde_ns_residual_timelevels_forward(ts, *pairs, nu, dt)
        ctx.ts, ctx.nu, ctx.dt = ts, nu, dt
        ctx.save_for_backward(*levels)
        return loss

    @staticmethod
    def backward(ctx, g):
        levels = ctx.saved_tensors
        pairs = [levels[i:i + 2] for i in range(0, 10, 2)]
        grads = _C.pde_ns_residual_timelevels_backward(ctx.ts, *pairs, ctx.nu, ctx.dt, g, tuple(ctx.needs_input_grad[3:13]))
        return (None,) * 3 + tuple(None if d is None else d.reshape(t.shape) for d, t in zip(grads, levels))


def navier_stokes_residual_loss_timelevels(time_samples, u, ux, wx, wxx, w, dt, nu):
    """NAVIER_STOKES with the time integration inside the kernel (model_pn.py:794-805, 818, 824, 843-845): every argument
    a pair (level [-1], level [-2]); w = vorticity(ux) of each level (wt = w[-1] - w[-2] is formed in the kernel)."""
    return _NsResidualLossLevels.apply(time_samples, float(nu), float(dt), u[0], u[1], ux[0], ux[1], wx[0], wx[1],
                                       wxx[0], wxx[1], w[0], w[1])
