"""Covariance / conic construction of kr4b/pigs' `gaussians.py:163-193`, name for name, as ONE fused CUDA kernel with
its own backward instead of the reference's chain of ~8 torch kernels and a batched
`torch.inverse` per training step:

    covariances, conics = build_covariances(scaling, transform)        # flattened [P, 3] ([xx, xy, yy]) in 2-D
    full_cov, full_conics = build_full_covariances(scaling, transform)  # [P, d, d]
    flatten_covariances(full_cov, full_conics)                          # [..., [0, 1, 3]] of the 2x2 matrices

A maintainer switches by replacing `gaussians.build_covariances` with `pigs_b200.gaussians.build_covariances`
(INTEGRATION.md).  No CPU path: the tensors must be CUDA fp32 (the reference chain is device-agnostic torch).
"""
import torch

from . import _C


class _BuildCovariances(torch.autograd.Function):
    @staticmethod
    def forward(ctx, scaling, transform):
        cov, con = _C.build_covariances(scaling, transform)
        ctx.save_for_backward(scaling, transform)
        ctx.set_materialize_grads(False)  # an unused output reaches the C ABI as a NULL upstream gradient
        return cov, con

    @staticmethod
    def backward(ctx, g_cov, g_con):
        scaling, transform = ctx.saved_tensors
        if g_cov is None and g_con is None:
            return None, None
        ds, dt = _C.build_covariances_backward(scaling, transform, g_cov, g_con)
        return ds.reshape(scaling.shape), dt.reshape(transform.shape)


def build_covariances(s, t):
    """gaussians.py:191-193: flattened covariances and conics, [P, d(d+1)/2] each (d = s.shape[-1], 1 or 2)."""
    return _BuildCovariances.apply(s, t)


def build_full_covariances(s, t):
    """gaussians.py:163-183: full [P, d, d] covariance and conic matrices."""
    cov, con = build_covariances(s, t)
    d = s.shape[-1]
    if d == 1:
        return cov.reshape(-1, 1, 1), con.reshape(-1, 1, 1)
    idx = torch.tensor([0, 1, 1, 2], device=cov.device)
    return cov[:, idx].reshape(-1, 2, 2), con[:, idx].reshape(-1, 2, 2)


def flatten_covariances(covariances, conics):
    """gaussians.py:185-189 (2-D only, like the reference)."""
    covariances = covariances.reshape(*covariances.shape[:-2], -1)[..., [0, 1, 3]]
    conics = conics.reshape(*conics.shape[:-2], -1)[..., [0, 1, 3]]
    return covariances, conics


# ---- dense all-pairs sampling (gaussians.py:48-58, 89-101, 103-116) ----------------
class _DenseSample(torch.autograd.Function):
    """One launch for any subset of (u, grad u, Hessian); backward = what torch.autograd gives the reference's dense
    functions (gradients to means, FULL conics and values; none to samples)."""

    @staticmethod
    def forward(ctx, mask, means, conics, values, samples):
        outs = _C.dense_forward(mask, means, conics, values, samples)
        ctx.mask = mask
        ctx.save_for_backward(means, conics, values, samples)
        ctx.shapes = (conics.shape, values.shape)
        return tuple(o for o in outs if o is not None)

    @staticmethod
    def backward(ctx, *grads):
        means, conics, values, samples = ctx.saved_tensors
        it = iter(grads)
        g = [next(it) if ctx.mask & bit else None for bit in (1, 2, 4)]
        mask = sum(bit for bit, x in zip((1, 2, 4), g) if x is not None)
        if mask == 0:
            return None, None, None, None, None
        dm, dc, dv = _C.dense_backward(mask, means, conics, values, samples, *g)
        return None, dm.reshape(means.shape), dc.reshape(ctx.shapes[0]), dv.reshape(ctx.shapes[1]), None


def dense_sample(mask, means, conics, values, samples):
    """(extension) any subset of u / grad u / Hessian of the dense path in ONE launch; mask bits 1 | 2 | 4.
    Returns a tuple in that order."""
    return _DenseSample.apply(int(mask), means, conics, values, samples)


def sample_gaussians(means, conics, values, samples):
    """gaussians.py:48-58: [N, C]; conics FULL [n, d, d]; samples [N, d] (any shape with N*d elements)."""
    return dense_sample(1, means, conics, values, samples)[0]


def gaussian_derivative(means, conics, values, samples):
    """gaussians.py:89-101: [N, d, C]."""
    return dense_sample(2, means, conics, values, samples)[0]


def gaussian_derivative2(means, conics, values, samples):
    """gaussians.py:103-116: [N, d, d, C]."""
    return dense_sample(4, means, conics, values, samples)[0]
