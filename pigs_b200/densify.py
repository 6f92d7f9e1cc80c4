"""Split / prune of the Gaussian set as a device-side compaction — the step that changes P
between training iterations and forces re-binning.  Mirrors, on CUDA fp32 tensors and with ONE host readback (the two
row counts):

    model_pn.py:578-610  Model.split(indices)      -> split(means, scaling, transforms, u, covariances, indices)
    model_pn.py:703-716  prune by keep_indices     -> plan(keep) + gather_rows
    test_no_mlp_1d.py:203-247 keep + clones, Adam  -> plan(keep, extend, m=1) + gather_rows(..., appended_scale=0)

`plan` returns the parent row of every output row (`src`), which is also what the optimiser-state surgery of
test_no_mlp_1d.py needs.  No CPU path."""
import ctypes

import torch

from . import _lib
from ._C import _bytes, _f32c, _ptr, _stream
from ._lib import check


def _mask(t, name, P):
    if t is None:
        return None
    if not t.is_cuda or t.numel() != P:
        raise RuntimeError(f"{name} must be a CUDA mask with one entry per row")
    return t.reshape(-1).to(torch.uint8).contiguous()  # bool -> u8 (a view-compatible copy)


def plan(P, keep=None, extend=None, m=1, device=None):
    """-> (src int64[P_out], n_keep, n_extend): output rows = kept rows in order, then m rows per extended row."""
    L = _lib.lib()
    ref = keep if keep is not None else extend
    dev = device if ref is None else ref.device
    keep8, ext8 = _mask(keep, "keep", P), _mask(extend, "extend", P)
    if P == 0:
        return torch.empty(0, dtype=torch.int64, device=dev), 0, 0
    ws = _bytes(L.pigs_compact_workspace_bytes(P), dev)
    st = _stream()
    check(L.pigs_compact_count(P, _ptr(keep8), _ptr(ext8), _ptr(ws), st), "pigs_compact_count")
    n_keep, n_ext = (int(v) for v in ws[:16].view(torch.int64).tolist())  # the one host readback
    src = torch.empty(n_keep + m * n_ext, dtype=torch.int64, device=dev)
    if src.numel():
        check(L.pigs_compact_fill(P, _ptr(keep8), _ptr(ext8), int(m), _ptr(ws), n_keep, _ptr(src), st), "pigs_compact_fill")
    return src, n_keep, n_ext


def _gather_rows_raw(x, src, n_keep, appended_scale):
    L = _lib.lib()
    x = _f32c(x, "x")
    W = x.numel() // max(x.shape[0], 1)
    out = torch.empty((src.numel(),) + tuple(x.shape[1:]), dtype=torch.float32, device=x.device)
    if src.numel():
        check(L.pigs_gather_rows(src.numel(), W, n_keep, _ptr(src), _ptr(x), _ptr(out), ctypes.c_float(appended_scale),
                                 _stream()), "pigs_gather_rows")
    return out


def _scatter_rows(g, src, n_keep, appended_scale, P):
    """Adjoint of the gather: dL/dx[p] = sum over output rows i with src[i] == p of scale_i * g[i]."""
    g = g.contiguous()
    if appended_scale != 1.0 and src.numel() > n_keep:
        g = g.clone()
        g[n_keep:] *= appended_scale
    dx = torch.zeros((P,) + tuple(g.shape[1:]), dtype=g.dtype, device=g.device)
    if src.numel():
        dx.index_add_(0, src, g)
    return dx


class _GatherRows(torch.autograd.Function):
    """x[src] * scale as ONE autograd node: in the reference the gathered tensors (means, scaling, transforms, u) are
    non-leaf results of the dynamics network (model_pn.py:684-687) and autograd flows through `x[indices]`, the `/ 2`
    and the `cat` of Model.split (model_pn.py:591-601); a raw kernel call would cut the network off the loss silently."""

    @staticmethod
    def forward(ctx, x, src, n_keep, appended_scale):
        ctx.save_for_backward(src)
        ctx.meta = (int(n_keep), float(appended_scale), x.shape[0])
        return _gather_rows_raw(x, src, n_keep, appended_scale)

    @staticmethod
    def backward(ctx, g):
        (src,) = ctx.saved_tensors
        n_keep, scale, P = ctx.meta
        return _scatter_rows(g, src, n_keep, scale, P), None, None, None


class _SplitMeans(torch.autograd.Function):
    """means of Model.split: kept rows copied, children at mu -/+ displacement.  The displacement comes out of a
    no_grad eigen-decomposition in the reference (model_pn.py:584-590), i.e. it is a constant: d out / d means[src] = I."""

    @staticmethod
    def forward(ctx, means, cov, src, n_keep):
        L = _lib.lib()
        D = means.shape[-1]
        ctx.save_for_backward(src)
        ctx.P = means.shape[0]
        out = torch.empty((src.numel(), D), dtype=torch.float32, device=means.device)
        if src.numel():
            check(L.pigs_split_means(src.numel(), D, n_keep, _ptr(src), _ptr(means), _ptr(cov), _ptr(out), _stream()),
                  "pigs_split_means")
        return out

    @staticmethod
    def backward(ctx, g):
        (src,) = ctx.saved_tensors
        return _scatter_rows(g, src, src.numel(), 1.0, ctx.P), None, None, None


def gather_rows(x, src, n_keep, appended_scale=1.0, appended_value=None):
    """x[src] with the appended rows (index >= n_keep) multiplied by `appended_scale` (0 -> exact zeros).  Differentiable
    with respect to x (the adjoint is the scatter-add by src with the same scale).
    Rows that are not float32 (boundary_mask is bool, ids are integers) take torch's own indexing; `appended_value`, when
    given, overwrites the appended rows — Model.split appends `True` mask rows and zero boundary rows rather than the
    parent's values (model_pn.py:602-605)."""
    if x.dtype != torch.float32:
        out = x[src]
        if appended_value is not None and src.numel() > n_keep:
            out[n_keep:] = appended_value
        return out
    out = _GatherRows.apply(_f32c(x, "x"), src, int(n_keep), float(appended_scale))
    if appended_value is not None and src.numel() > n_keep:
        out = out.clone()
        out[n_keep:] = appended_value
    return out


def split(means, scaling, transforms, u, covariances, indices):
    """Model.split (model_pn.py:578-610): rows not in `indices` are kept, every row in `indices` is replaced by two
    children at mu -/+ lambda_max * v_max of its covariance with the same scaling / transform and u / 2.
    covariances: flattened [P, d(d+1)/2] (or the full [P, d, d]).  Returns (means, scaling, transforms, u, src).
    Gradients flow to means, scaling, transforms and u exactly as through the reference's indexing / cat / `/ 2`; the
    displacement is a constant (the reference computes it under no_grad)."""
    means = _f32c(means, "means")
    P, D = means.shape[0], means.shape[-1]
    cov = _f32c(covariances.detach(), "covariances")
    if cov.dim() == 3:  # full matrices: flatten like gaussians.py:186-189
        cov = cov.reshape(P, -1)[:, [0, 1, 3]].contiguous() if D == 2 else cov.reshape(P, 1)
    idx = indices.reshape(-1).bool()
    src, n_keep, n_split = plan(P, keep=~idx, extend=idx, m=2)
    out_means = _SplitMeans.apply(means, cov, src, n_keep)
    return (out_means, gather_rows(scaling, src, n_keep), gather_rows(transforms, src, n_keep),
            gather_rows(u, src, n_keep, appended_scale=0.5), src)
