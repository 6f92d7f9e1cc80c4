"""The reference's 12 native entry points (diff-gaussian-sampling/ext.cpp:20-31), same names, same
argument order, same return values — implemented as thin tensor -> pointer glue over the C ABI of
libpigs_b200.so (include/pigs_b200.h).  torch owns every allocation; the library only launches
kernels on torch's current CUDA stream.

`debug=True` has the reference's meaning (cuda_sampler/auxiliary.h:33-40): the library synchronises the device and
checks for errors after EVERY kernel launch (pigs_set_debug), so a faulting kernel raises RuntimeError naming its own
launch site.

Differences a caller can observe (all documented in DESIGN.md):
  * the byte buffers (`binning_buffer`, `sample_binning_buffer`) have our own internal layout — they
    are opaque in the reference too (sample_points.cu:48-55);
  * preprocess costs ONE device->host readback (reference: two), forward/backward cost none
    (reference: one `.item()` each, sample_points.cu:124,174).
"""
import ctypes

import torch

from . import _lib
from ._lib import HEADER_BYTES, PigsHeader, check

FN_VALUE, FN_DERIV, FN_HESS, FN_THIRD = 1, 2, 4, 8
FN_LAPL = 16  # extension (include/pigs_b200.h PIGS_FN_LAPL): trace of the Hessian, [N, C]; excludes FN_HESS

_pinned = {}


def _ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _f32c(t, name):
    """fp32 + CUDA asserted; non-contiguous inputs are copied like the reference's `.contiguous()`."""
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name} must be a torch.Tensor")
    if t.dtype != torch.float32:
        raise RuntimeError(f"{name}: expected scalar type Float but found {t.dtype}")  # reference: .data<float>() throws
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (pigs_b200 has no CPU path)")
    return t.contiguous()


def _on_device(fn):
    """Run a binding function with the device of its first tensor argument current (the library launches on the current
    device's stream); the compiled binding does the same with a CUDAGuard."""
    import functools

    @functools.wraps(fn)
    def wrapped(*a, **k):
        t = next((x for x in a if isinstance(x, torch.Tensor) and x.is_cuda), None)
        if t is None or t.device.index == torch.cuda.current_device():
            return fn(*a, **k)
        with torch.cuda.device(t.device):
            return fn(*a, **k)
    return wrapped


def _bytes(n, device):
    return torch.empty(int(n), dtype=torch.uint8, device=device)


def _dims(means, values, samples):
    P = means.size(0)
    D = means.size(-1)
    N = samples.size(0)
    C = values.size(-1)
    return P, D, N, C


@_on_device
def sample_bounds(samples, D):
    """float32[2*D] on the device: per-dimension min then max of the samples (sample_points.cu:70-71)."""
    L = _lib.lib()
    samples = _f32c(samples, "samples")
    mm = torch.empty(2 * D, dtype=torch.float32, device=samples.device)
    check(L.pigs_sample_bounds(_ptr(samples), samples.numel() // D, D, _ptr(mm), _stream()), "pigs_sample_bounds")
    return mm


def read_header(buf):
    """One pinned-memory readback of the 256-byte device header at the start of `buf`."""
    dev = buf.device
    key = (dev.index, torch.cuda.current_stream().cuda_stream)
    host = _pinned.get(key)
    if host is None:
        host = torch.empty(HEADER_BYTES, dtype=torch.uint8).pin_memory()
        _pinned[key] = host
    host.copy_(buf[:HEADER_BYTES], non_blocking=True)
    torch.cuda.current_stream().synchronize()
    return PigsHeader.from_buffer_copy(host.numpy().tobytes())


def _read_i64(dev_scalar):
    """One device int64 -> host through pinned memory on the current stream (the only sync of the call)."""
    key = ("i64", dev_scalar.device.index, torch.cuda.current_stream().cuda_stream)
    host = _pinned.get(key)
    if host is None:
        host = torch.empty(1, dtype=torch.int64).pin_memory()
        _pinned[key] = host
    host.copy_(dev_scalar, non_blocking=True)
    torch.cuda.current_stream().synchronize()
    return int(host.item())


@_on_device
def preprocess_gaussians(means, values, covariances, conics, samples, debug, bounds=None, cull=None):
    """PreprocessCUDA (sample_points.cu:37-98).  Returns
    (num_rendered, binning_buffer, sample_binning_buffer, ranges, sample_ranges, radii).
    `bounds` (device float32[2*D], optional, extension) overrides the sample bounding box — used by
    the multi-GPU sampler so every rank bins on the global tile grid.
    `cull` = (sigmas, subdiv) (optional, extension): the opt-in culled mode of include/pigs_b200.h
    pigs_preprocess_count_culled — a SUBSET of the reference's pair set, not the reference's result."""
    L = _lib.lib()
    L.pigs_set_debug(1 if debug else 0)
    means = _f32c(means, "means")
    P, D, N, C = _dims(means, values, samples)
    dev = means.device
    radii = torch.zeros(P, dtype=torch.float32, device=dev)
    empty = torch.empty(0, dtype=torch.uint8, device=dev)
    if P == 0 or N == 0:
        return 0, empty, empty.clone(), empty.clone(), empty.clone(), radii
    if D not in (1, 2):
        raise RuntimeError(f"pigs_b200 supports 1-D and 2-D mixtures (got D={D}), like the reference kernels")
    cov = _f32c(covariances, "covariances")
    samples = _f32c(samples, "samples")
    _f32c(values, "values")
    if cov.numel() != P * D * (D + 1) // 2:
        raise RuntimeError("covariances must hold P*D*(D+1)/2 elements")
    st = _stream()
    mm = torch.empty(2 * D, dtype=torch.float32, device=dev) if bounds is None else _f32c(bounds, "bounds")
    geom = _bytes(L.pigs_geom_workspace_bytes(P), dev)
    # bounds + tile grid + radii / tiles touched + scan: one single-CTA launch for small problems, the grid-wide chain else
    subdiv, sigmas = (1, 0.0) if cull is None else (int(cull[1]), float(cull[0]))
    check(L.pigs_preprocess_count_fused(P, D, N, _ptr(means), _ptr(cov), _ptr(samples), _ptr(mm), 0 if bounds is None else 1,
                                        _ptr(radii), _ptr(geom), subdiv, sigmas, st), "pigs_preprocess_count_fused")
    h = read_header(geom)
    blocks, R = int(h.blocks), int(h.num_rendered)
    binning = _bytes(L.pigs_binning_bytes(R, blocks), dev)
    sbinning = _bytes(L.pigs_sample_binning_bytes(N, blocks, D), dev)
    ranges = torch.empty(blocks * 8 + 8, dtype=torch.uint8, device=dev)  # written (or zeroed) by the library
    sample_ranges = torch.empty(blocks * 8 + 8, dtype=torch.uint8, device=dev)
    sort_ws = _bytes(L.pigs_sort_workspace_bytes(R, N), dev)
    conics = _f32c(conics, "conics")
    check(L.pigs_preprocess_bin(P, D, N, blocks, R, _ptr(means), _ptr(conics), _ptr(samples), _ptr(radii), _ptr(geom), _ptr(binning),
                                _ptr(sbinning), _ptr(ranges), _ptr(sample_ranges), _ptr(sort_ws), st),
          "pigs_preprocess_bin")
    return R, binning, sbinning, ranges, sample_ranges, radii


def _blocks_of(ranges):
    return ranges.numel() // 8 - 1


@_on_device
def sample_forward(fn_mask, means, values, conics, samples, num_rendered, binning_buffer, sample_binning_buffer, ranges,
                   sample_ranges, debug):
    """Fused forward for any subset of {value, derivative, Hessian | Laplacian, third}.  Returns a 4-tuple with
    None for the functions not requested; with FN_LAPL slot 2 holds the Laplacian [N, C] instead of the Hessian."""
    L = _lib.lib()
    L.pigs_set_debug(1 if debug else 0)
    means = _f32c(means, "means")
    values = _f32c(values, "values")
    conics = _f32c(conics, "conics")
    samples = _f32c(samples, "samples")
    P, D, N, C = _dims(means, values, samples)
    dev = means.device
    if (fn_mask & FN_LAPL) and (fn_mask & FN_HESS):
        raise RuntimeError("FN_LAPL and FN_HESS are mutually exclusive (they share an output slot)")
    shp = lambda k: (N,) + (D,) * k + (C,)
    outs = [torch.zeros(shp(k), dtype=torch.float32, device=dev) if fn_mask & (1 << k) else None for k in range(4)]
    if fn_mask & FN_LAPL:
        outs[2] = torch.zeros((N, C), dtype=torch.float32, device=dev)
    if P != 0 and N != 0:
        blocks = _blocks_of(ranges)
        ws_bytes = L.pigs_eval_workspace_bytes(num_rendered, N, blocks, D, C, fn_mask)
        ws = _bytes(ws_bytes, dev)
        check(L.pigs_sample_forward(fn_mask, P, D, N, C, blocks, num_rendered, _ptr(means), _ptr(values), _ptr(conics),
                                    _ptr(samples), _ptr(binning_buffer), _ptr(sample_binning_buffer), _ptr(ranges),
                                    _ptr(sample_ranges), _ptr(outs[0]), _ptr(outs[1]), _ptr(outs[2]), _ptr(outs[3]),
                                    _ptr(ws), ws_bytes, _stream()), "pigs_sample_forward")
    return tuple(outs)


@_on_device
def sample_backward(fn_mask, means, values, conics, samples, num_rendered, grads, binning_buffer, sample_binning_buffer,
                    ranges, sample_ranges, debug):
    """Fused backward; `grads` is a 4-tuple of upstream gradients (None = zero; with FN_LAPL slot 2 is the
    Laplacian's [N, C] gradient).  Returns (dL_dmeans [P,D], dL_dvalues [P,C], dL_dconics [P,D(D+1)/2])."""
    L = _lib.lib()
    L.pigs_set_debug(1 if debug else 0)
    means = _f32c(means, "means")
    values = _f32c(values, "values")
    conics = _f32c(conics, "conics")
    samples = _f32c(samples, "samples")
    P, D, N, C = _dims(means, values, samples)
    dev = means.device
    # one flat buffer, three contiguous views: a multi-GPU caller all-reduces `flat` with one collective
    K = D * (D + 1) // 2
    flat = torch.zeros(P * (D + C + K), dtype=torch.float32, device=dev)
    dm = flat[:P * D].view(P, D)
    dv = flat[P * D:P * (D + C)].view(P, C)
    dc = flat[P * (D + C):].view(P, K)
    g = [None if x is None else _f32c(x, "dL_dout") for x in grads]
    if P != 0 and N != 0:
        blocks = _blocks_of(ranges)
        ws_bytes = L.pigs_eval_workspace_bytes(num_rendered, N, blocks, D, C, fn_mask)
        ws = _bytes(ws_bytes, dev)
        check(L.pigs_sample_backward(fn_mask, P, D, N, C, blocks, num_rendered, _ptr(means), _ptr(values), _ptr(conics),
                                     _ptr(samples), _ptr(binning_buffer), _ptr(sample_binning_buffer), _ptr(ranges),
                                     _ptr(sample_ranges), _ptr(g[0]), _ptr(g[1]), _ptr(g[2]), _ptr(g[3]), _ptr(dm),
                                     _ptr(dv), _ptr(dc), _ptr(ws), ws_bytes, _stream()), "pigs_sample_backward")
    return dm, dv, dc


def packed_base(dm):
    """The flat [P*(D+C+K)] gradient buffer that sample_backward's three results are views of."""
    base = dm._base
    if base is None:
        raise RuntimeError("not a view returned by sample_backward")
    return base


def _single(bit):
    k = bit.bit_length() - 1

    def fwd(means, values, conics, samples, num_rendered, binning_buffer, sample_binning_buffer, ranges, sample_ranges,
            debug):
        return sample_forward(bit, means, values, conics, samples, num_rendered, binning_buffer, sample_binning_buffer,
                              ranges, sample_ranges, debug)[k]

    def bwd(means, values, conics, samples, num_rendered, dL_dout, binning_buffer, sample_binning_buffer, ranges,
            sample_ranges, debug):
        grads = [None] * 4
        grads[k] = dL_dout
        return sample_backward(bit, means, values, conics, samples, num_rendered, grads, binning_buffer,
                               sample_binning_buffer, ranges, sample_ranges, debug)

    return fwd, bwd


# ext.cpp:21-28
sample_gaussians, sample_gaussians_backward = _single(FN_VALUE)
sample_gaussians_derivative, sample_gaussians_derivative_backward = _single(FN_DERIV)
sample_gaussians_laplacian, sample_gaussians_laplacian_backward = _single(FN_HESS)
sample_gaussians_third_derivative, sample_gaussians_third_derivative_backward = _single(FN_THIRD)


def point_lists(num_rendered, binning_buffer, sample_binning_buffer, N):
    """(point_list u32[R], point_tiles u32[R], sample_point_list u32[N]) as int32 tensor views of the
    opaque buffers — for parity tests / debugging (the reference's BinningState::point_list)."""
    L = _lib.lib()

    def view(buf, addr, n):
        off = addr - buf.data_ptr()
        return buf[off:off + 4 * n].view(torch.int32)

    pl = view(binning_buffer, L.pigs_point_list(_ptr(binning_buffer)), num_rendered)
    pt = view(binning_buffer, L.pigs_point_tiles(_ptr(binning_buffer), num_rendered), num_rendered)
    sl = view(sample_binning_buffer, L.pigs_sample_list(_ptr(sample_binning_buffer)), N)
    return pl, pt, sl


# ---- neighbour aggregation (ext.cpp:29-31) ----------------------------------------------------------
@_on_device
def preprocess_aggregate(means, conics, radii, debug, rows=None):
    """AggregateNeighborsPreprocessCUDA (aggregate_neighbors.cu:323-367).  Returns
    (indices i64[M], ranges i64[P], dists f32[M,D], densities f32[M], inv_total_densities f32[P]).
    rows = (lo, hi) (extension, multi-GPU): lists of the query Gaussians [lo, hi) only; the others get empty rows."""
    L = _lib.lib()
    L.pigs_set_debug(1 if debug else 0)
    means = _f32c(means, "means")
    conics = _f32c(conics, "conics")
    radii = _f32c(radii, "radii")
    P, D = means.size(0), means.size(-1)
    lo, hi = (0, P) if rows is None else (int(rows[0]), int(rows[1]))
    dev = means.device
    ws = _bytes(max(int(L.pigs_neighbor_workspace_bytes(P, D)), 16), dev)
    counts = torch.zeros(P, dtype=torch.int64, device=dev)
    st = _stream()
    check(L.pigs_neighbors_count_rows(P, D, _ptr(means), _ptr(radii), lo, hi, _ptr(counts), _ptr(ws), st),
          "pigs_neighbors_count_rows")
    # CSR ranges = inclusive scan of the counts (library scan), total read back once through pinned memory
    ranges = torch.empty_like(counts)
    total = torch.zeros(1, dtype=torch.int64, device=dev)
    length = 0
    if P > 0:
        sws = _bytes(max(int(L.pigs_neighbors_scan_workspace_bytes(P)), 16), dev)
        check(L.pigs_neighbors_ranges(P, _ptr(counts), _ptr(ranges), _ptr(total), _ptr(sws), st), "pigs_neighbors_ranges")
        length = _read_i64(total)
    # every CSR slot is written by the fill kernels (skipped neighbours get index -1 / density 0 there)
    indices = torch.empty((length,), dtype=torch.int64, device=dev)
    dists = torch.empty((length, D), dtype=torch.float32, device=dev)
    densities = torch.empty((length,), dtype=torch.float32, device=dev)
    inv_total = torch.zeros((P,), dtype=torch.float32, device=dev)
    check(L.pigs_neighbors_fill_rows(P, D, _ptr(means), _ptr(conics), _ptr(radii), lo, hi, _ptr(ranges), _ptr(indices),
                                     _ptr(dists), _ptr(densities), _ptr(inv_total), _ptr(ws), st), "pigs_neighbors_fill_rows")
    return indices, ranges, dists, densities, inv_total


def _f64c(t, name):
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (pigs_b200 has no CPU path)")
    return t.to(torch.float64).contiguous()


def _aggregate_neighbors_f64(features, transform, queries, keys, frequencies, distance_transform, indices, ranges, dists,
                             densities, inv_total_densities, debug):
    """float64 path = the reference built with `typedef double FLOAT` (config.h:20), for gradcheck
    (test_neighbor_aggregation.py:94-101): generic kernels instantiated for double, CSR values widened here."""
    L_ = _lib.lib()
    L_.pigs_set_debug(1 if debug else 0)
    a = [_f64c(t, n) for t, n in ((features, "features"), (transform, "transform"), (queries, "queries"), (keys, "keys"),
                                  (frequencies, "frequencies"), (distance_transform, "distance_transform"))]
    dists, densities, inv_total = _f64c(dists, "dists"), _f64c(densities, "densities"), _f64c(inv_total_densities, "inv_total")
    P, D, L = a[0].size(0), dists.size(-1), a[0].size(-1)
    K, E = a[2].size(-1), a[5].size(-1) // 2
    w, e, f = torch.zeros_like(densities), torch.zeros_like(densities), torch.zeros_like(densities)
    out = torch.zeros((P, L), dtype=torch.float64, device=a[0].device)
    if P > 0:
        check(L_.pigs_aggregate_forward_f64(P, D, L, K, E, *[_ptr(t) for t in a], _ptr(indices.contiguous()),
                                            _ptr(ranges.contiguous()), _ptr(dists), _ptr(densities), _ptr(inv_total), _ptr(w),
                                            _ptr(e), _ptr(f), _ptr(out), _stream()), "pigs_aggregate_forward_f64")
    return w, e, f, out


def _aggregate_neighbors_backward_f64(features, transform, queries, keys, frequencies, distance_transform, indices, ranges,
                                      dists, densities, weights, embeddings, factors, inv_total_densities, g, debug):
    L_ = _lib.lib()
    L_.pigs_set_debug(1 if debug else 0)
    a = [_f64c(t, n) for t, n in ((features, "features"), (transform, "transform"), (queries, "queries"), (keys, "keys"),
                                  (frequencies, "frequencies"), (distance_transform, "distance_transform"))]
    dists, densities, inv_total = _f64c(dists, "dists"), _f64c(densities, "densities"), _f64c(inv_total_densities, "inv_total")
    saved = [_f64c(t, "saved") for t in (weights, embeddings, factors)]
    g = _f64c(g, "dL_dneighbor_features")
    P, D, L = a[0].size(0), dists.size(-1), a[0].size(-1)
    K, E = a[2].size(-1), a[5].size(-1) // 2
    outs = [torch.zeros_like(t) for t in a]
    if P > 0:
        check(L_.pigs_aggregate_backward_f64(P, D, L, K, E, *[_ptr(t) for t in a], _ptr(indices.contiguous()),
                                             _ptr(ranges.contiguous()), _ptr(dists), _ptr(densities), *[_ptr(t) for t in saved],
                                             _ptr(inv_total), _ptr(g), *[_ptr(t) for t in outs], _stream()),
              "pigs_aggregate_backward_f64")
    return tuple(outs)


@_on_device
def aggregate_neighbors(features, transform, queries, keys, frequencies, distance_transform, indices, ranges, dists,
                        densities, inv_total_densities, debug):
    """AggregateNeighborsCUDA (aggregate_neighbors.cu:369-415).  Returns
    (weights, embeddings, factors [M], neighbor_features [P,L])."""
    if features.dtype == torch.float64:
        return _aggregate_neighbors_f64(features, transform, queries, keys, frequencies, distance_transform, indices, ranges,
                                        dists, densities, inv_total_densities, debug)
    L_ = _lib.lib()
    L_.pigs_set_debug(1 if debug else 0)
    features = _f32c(features, "features")
    transform = _f32c(transform, "transform")
    queries = _f32c(queries, "queries")
    keys = _f32c(keys, "keys")
    frequencies = _f32c(frequencies, "frequencies")
    dt = _f32c(distance_transform, "distance_transform")
    dists = _f32c(dists, "dists")
    densities = _f32c(densities, "densities")
    inv_total = _f32c(inv_total_densities, "inv_total_densities")
    indices = indices.contiguous()
    ranges = ranges.contiguous()
    P, D, L = features.size(0), dists.size(-1), features.size(-1)
    K, E = queries.size(-1), dt.size(-1) // 2
    dev = features.device
    # no zero fills: the kernels write every slot of every row (skipped neighbours as zeros) and every output row
    weights = torch.empty_like(densities)
    embeddings = torch.empty_like(densities)
    factors = torch.empty_like(densities)
    out = torch.empty((P, L), dtype=torch.float32, device=dev)
    if P > 0:
        check(L_.pigs_aggregate_forward(P, D, L, K, E, _ptr(features), _ptr(transform), _ptr(queries), _ptr(keys),
                                        _ptr(frequencies), _ptr(dt), _ptr(indices), _ptr(ranges), _ptr(dists),
                                        _ptr(densities), _ptr(inv_total), _ptr(weights), _ptr(embeddings), _ptr(factors),
                                        _ptr(out), _stream()), "pigs_aggregate_forward")
    return weights, embeddings, factors, out


@_on_device
def aggregate_neighbors_backward(features, transform, queries, keys, frequencies, distance_transform, indices, ranges,
                                 dists, densities, weights, embeddings, factors, inv_total_densities,
                                 dL_dneighbor_features, debug):
    """AggregateNeighborsBackwardCUDA (aggregate_neighbors.cu:417-475)."""
    if features.dtype == torch.float64:
        return _aggregate_neighbors_backward_f64(features, transform, queries, keys, frequencies, distance_transform, indices,
                                                 ranges, dists, densities, weights, embeddings, factors, inv_total_densities,
                                                 dL_dneighbor_features, debug)
    L_ = _lib.lib()
    L_.pigs_set_debug(1 if debug else 0)
    features = _f32c(features, "features")
    transform = _f32c(transform, "transform")
    queries = _f32c(queries, "queries")
    keys = _f32c(keys, "keys")
    frequencies = _f32c(frequencies, "frequencies")
    dt = _f32c(distance_transform, "distance_transform")
    dists = _f32c(dists, "dists")
    g = _f32c(dL_dneighbor_features, "dL_dneighbor_features")
    P, D, L = features.size(0), dists.size(-1), features.size(-1)
    K, E = queries.size(-1), dt.size(-1) // 2
    M = densities.numel()
    dev = features.device
    gf = torch.zeros_like(features)
    gt = torch.zeros_like(transform)
    gq = torch.zeros_like(queries)
    gk = torch.zeros_like(keys)
    gfr = torch.zeros_like(frequencies)
    gdt = torch.zeros_like(dt)
    if P > 0:
        ws = _bytes(max(int(L_.pigs_aggregate_backward_workspace_bytes(P, L, K, E)), 16), dev)
        check(L_.pigs_aggregate_backward(P, D, L, K, E, M, _ptr(features), _ptr(transform), _ptr(queries), _ptr(keys),
                                         _ptr(frequencies), _ptr(dt), _ptr(indices.contiguous()), _ptr(ranges.contiguous()),
                                         _ptr(dists), _ptr(_f32c(densities, "densities")), _ptr(_f32c(weights, "weights")),
                                         _ptr(_f32c(embeddings, "embeddings")), _ptr(_f32c(factors, "factors")),
                                         _ptr(_f32c(inv_total_densities, "inv_total_densities")), _ptr(g), _ptr(gf),
                                         _ptr(gt), _ptr(gq), _ptr(gk), _ptr(gfr), _ptr(gdt), _ptr(ws), _stream()),
              "pigs_aggregate_backward")
    return gf, gt, gq, gk, gfr, gdt


def _disp(t, rows, width):
    """Base pointer of a [hi-lo, width] window tensor displaced by -lo rows: the `_rows` entry points address the
    row-indexed arrays with the GLOBAL row index and touch rows [lo, hi) only (include/pigs_b200.h)."""
    return ctypes.c_void_p(t.data_ptr() - int(rows[0]) * int(width) * 4)


@_on_device
def aggregate_neighbors_rows(features, transform, queries_rows, keys, frequencies, distance_transform, indices, ranges,
                             dists, densities, inv_total_densities, rows, debug):
    """(extension, multi-GPU) aggregate_neighbors for the query rows [lo, hi) = `rows` only.  features / keys are
    full-height [P, .] buffers of which the rows the window's CSR names must be valid; queries_rows is the window's
    [hi-lo, K]; ranges / inv_total_densities are the [P] arrays of preprocess_aggregate(rows=...).
    Returns (weights, embeddings, factors [M_local], neighbor_features [hi-lo, L])."""
    L_ = _lib.lib()
    L_.pigs_set_debug(1 if debug else 0)
    features = _f32c(features, "features")
    transform = _f32c(transform, "transform")
    queries_rows = _f32c(queries_rows, "queries")
    keys = _f32c(keys, "keys")
    frequencies = _f32c(frequencies, "frequencies")
    dt = _f32c(distance_transform, "distance_transform")
    dists = _f32c(dists, "dists")
    densities = _f32c(densities, "densities")
    inv_total = _f32c(inv_total_densities, "inv_total_densities")
    indices, ranges = indices.contiguous(), ranges.contiguous()
    lo, hi = int(rows[0]), int(rows[1])
    P, D, L = features.size(0), dists.size(-1), features.size(-1)
    K, E = queries_rows.size(-1), dt.size(-1) // 2
    if queries_rows.size(0) != hi - lo or keys.size(0) != P or ranges.numel() != P:
        raise RuntimeError("aggregate_neighbors_rows: queries must hold the window's rows, features / keys / ranges all P")
    weights = torch.empty_like(densities)
    embeddings = torch.empty_like(densities)
    factors = torch.empty_like(densities)
    out = torch.empty((hi - lo, L), dtype=torch.float32, device=features.device)
    if hi > lo:
        check(L_.pigs_aggregate_forward_rows(P, D, L, K, E, lo, hi, _ptr(features), _ptr(transform),
                                             _disp(queries_rows, rows, K), _ptr(keys), _ptr(frequencies), _ptr(dt),
                                             _ptr(indices), _ptr(ranges), _ptr(dists), _ptr(densities), _ptr(inv_total),
                                             _ptr(weights), _ptr(embeddings), _ptr(factors), _disp(out, rows, L), _stream()),
              "pigs_aggregate_forward_rows")
    return weights, embeddings, factors, out


@_on_device
def aggregate_neighbors_rows_backward(features, transform, queries_rows, keys, frequencies, distance_transform, indices,
                                      ranges, dists, densities, weights, embeddings, factors, inv_total_densities,
                                      dL_dout_rows, g_features, g_keys, rows, debug):
    """Backward of aggregate_neighbors_rows.  g_features / g_keys are caller-provided full-height [P, .] buffers the
    kernel ADDS this window's contributions into (rows named by the window's CSR slots); returns
    (dL_dtransform, dL_dqueries [hi-lo, K], dL_dfrequencies, dL_ddistance_transform) of this window."""
    L_ = _lib.lib()
    L_.pigs_set_debug(1 if debug else 0)
    features = _f32c(features, "features")
    transform = _f32c(transform, "transform")
    queries_rows = _f32c(queries_rows, "queries")
    keys = _f32c(keys, "keys")
    frequencies = _f32c(frequencies, "frequencies")
    dt = _f32c(distance_transform, "distance_transform")
    dists = _f32c(dists, "dists")
    g = _f32c(dL_dout_rows, "dL_dneighbor_features")
    lo, hi = int(rows[0]), int(rows[1])
    P, D, L = features.size(0), dists.size(-1), features.size(-1)
    K, E = queries_rows.size(-1), dt.size(-1) // 2
    if g.size(0) != hi - lo or not (g_features.is_contiguous() and g_keys.is_contiguous()):
        raise RuntimeError("aggregate_neighbors_rows_backward: dL_dout must hold the window's rows; gradient buffers contiguous")
    gt = torch.zeros_like(transform)
    gq = torch.zeros_like(queries_rows)
    gfr = torch.zeros_like(frequencies)
    gdt = torch.zeros_like(dt)
    if hi > lo:
        check(L_.pigs_aggregate_backward_rows(P, D, L, K, E, densities.numel(), lo, hi, _ptr(features), _ptr(transform),
                                              _disp(queries_rows, rows, K), _ptr(keys), _ptr(frequencies), _ptr(dt),
                                              _ptr(indices.contiguous()), _ptr(ranges.contiguous()), _ptr(dists),
                                              _ptr(_f32c(densities, "densities")), _ptr(_f32c(weights, "weights")),
                                              _ptr(_f32c(embeddings, "embeddings")), _ptr(_f32c(factors, "factors")),
                                              _ptr(_f32c(inv_total_densities, "inv_total_densities")), _disp(g, rows, L),
                                              _ptr(g_features), _ptr(gt), _disp(gq, rows, K), _ptr(g_keys), _ptr(gfr),
                                              _ptr(gdt), None, _stream()), "pigs_aggregate_backward_rows")
    return gt, gq, gfr, gdt


def build_covariances(scaling, transform):
    """(scaling [P, D], transform [P, 1]) -> flattened (covariances, conics) [P, D(D+1)/2]; gaussians.py:163-193."""
    L = _lib.lib()
    scaling = _f32c(scaling, "scaling")
    P, D = scaling.shape[0], scaling.shape[-1]
    transform = _f32c(transform, "transform") if D > 1 else None
    n = D * (D + 1) // 2
    cov = torch.empty((P, n), dtype=torch.float32, device=scaling.device)
    con = torch.empty((P, n), dtype=torch.float32, device=scaling.device)
    check(L.pigs_build_covariances_forward(P, D, _ptr(scaling), _ptr(transform), _ptr(cov), _ptr(con), _stream()),
          "pigs_build_covariances_forward")
    return cov, con


def build_covariances_backward(scaling, transform, g_cov, g_con):
    L = _lib.lib()
    scaling = _f32c(scaling, "scaling")
    P, D = scaling.shape[0], scaling.shape[-1]
    transform = _f32c(transform, "transform") if D > 1 else None
    g_cov = None if g_cov is None else _f32c(g_cov, "dL_dcovariances")
    g_con = None if g_con is None else _f32c(g_con, "dL_dconics")
    ds = torch.empty_like(scaling)
    dt = torch.empty((P, 1), dtype=torch.float32, device=scaling.device)
    check(L.pigs_build_covariances_backward(P, D, _ptr(scaling), _ptr(transform), _ptr(g_cov), _ptr(g_con), _ptr(ds),
                                            _ptr(dt), _stream()), "pigs_build_covariances_backward")
    return ds, dt


# ---- PDE right-hand side + residual loss (model_pn.py:612-625, 822-834) ------------
PDE_DIFFUSION, PDE_BURGERS, PDE_WAVE = 1, 3, 4  # model_pn.py:13-19 Problem enum values


def _pde_args(problem, u, ux, lap, ut):
    lap = _f32c(lap, "lap")
    ut = _f32c(ut, "ut")
    N = lap.shape[0]
    C = lap.numel() // max(N, 1)
    if ut.numel() != lap.numel():
        raise RuntimeError("ut and lap must hold N*C elements each")
    u = None if u is None else _f32c(u, "u")
    ux = None if ux is None else _f32c(ux, "ux")
    D = 2 if ux is None else ux.numel() // max(N * C, 1)
    if problem in (PDE_BURGERS, PDE_WAVE) and (u is None or u.numel() != lap.numel()):
        raise RuntimeError("u [N, C] is required for this problem")
    if problem == PDE_BURGERS and (ux is None or D not in (1, 2)):
        raise RuntimeError("ux [N, D, C] is required for BURGERS")
    return N, D if D in (1, 2) else 2, C, u, ux, lap, ut


def pde_residual_forward(problem, u, ux, lap, ut, nu, dt, want_residual=False):
    """-> (loss: 0-dim device tensor, residual [N, C] or None)."""
    L = _lib.lib()
    N, D, C, u, ux, lap, ut = _pde_args(problem, u, ux, lap, ut)
    dev = lap.device
    loss = torch.zeros((), dtype=torch.float32, device=dev)
    res = torch.empty_like(lap) if want_residual else None
    if N > 0:
        ws = _bytes(max(int(L.pigs_pde_residual_workspace_bytes(N, C)), 16), dev)
        check(L.pigs_pde_residual_forward(problem, N, D, C, float(nu), float(dt), _ptr(u), _ptr(ux), _ptr(lap), _ptr(ut),
                                          _ptr(res), _ptr(loss), _ptr(ws), _stream()), "pigs_pde_residual_forward")
    return loss, res


def pde_residual_backward(problem, u, ux, lap, ut, nu, dt, g_loss, need=(True, True, True, True)):
    """-> (dL_du, dL_dux, dL_dlap, dL_dut), None where not needed or identically absent (DIFFUSION: u, ux)."""
    L = _lib.lib()
    N, D, C, u, ux, lap, ut = _pde_args(problem, u, ux, lap, ut)
    g_loss = _f32c(g_loss, "dL_dloss").reshape(1)
    uses_u = problem in (PDE_BURGERS, PDE_WAVE)
    d_u = torch.empty_like(u) if (need[0] and uses_u) else None
    d_ux = torch.empty_like(ux) if (need[1] and problem == PDE_BURGERS) else None
    d_lap = torch.empty_like(lap) if need[2] else None
    d_ut = torch.empty_like(ut) if need[3] else None
    if N > 0:
        check(L.pigs_pde_residual_backward(problem, N, D, C, float(nu), float(dt), _ptr(u), _ptr(ux), _ptr(lap), _ptr(ut),
                                           _ptr(g_loss), _ptr(d_u), _ptr(d_ux), _ptr(d_lap), _ptr(d_ut), _stream()),
              "pigs_pde_residual_backward")
    return d_u, d_ux, d_lap, d_ut


# ---- NAVIER_STOKES residual (vorticity form) + vorticity extraction (model_pn.py:629-631, 779-781, 818, 843-845) ----
def _al16(t):
    """the NAVIER_STOKES / vorticity kernels read whole rows as 8-/16-byte vectors: a view at an odd float offset is copied"""
    return t if t.data_ptr() % 16 == 0 else t.clone(memory_format=torch.contiguous_format)


def _ns_args(u, ux, wx, wxx, wt):
    u, ux, wx, wxx, wt = (_al16(_f32c(t, n)) for t, n in ((u, "u"), (ux, "ux"), (wx, "wx"), (wxx, "wxx"), (wt, "wt")))
    N = wt.numel()
    if u.numel() != 2 * N or ux.numel() != 4 * N or wx.numel() != 2 * N or wxx.numel() != 4 * N:
        raise RuntimeError("NAVIER_STOKES residual: u [N,2], ux [N,2,2], wx [N,2], wxx [N,2,2], wt [N] expected")
    return N, u, ux, wx, wxx, wt


def pde_ns_residual_forward(u, ux, wx, wxx, wt, nu, dt, want_residual=False):
    """-> (loss: 0-dim device tensor, residual [N] or None, divergence [N] or None)."""
    L = _lib.lib()
    N, u, ux, wx, wxx, wt = _ns_args(u, ux, wx, wxx, wt)
    dev = wt.device
    loss = torch.zeros((), dtype=torch.float32, device=dev)
    res = torch.empty(N, dtype=torch.float32, device=dev) if want_residual else None
    div = torch.empty(N, dtype=torch.float32, device=dev) if want_residual else None
    if N > 0:
        ws = _bytes(max(int(L.pigs_pde_residual_workspace_bytes(N, 2)), 16), dev)
        check(L.pigs_pde_ns_residual_forward(N, float(nu), float(dt), _ptr(u), _ptr(ux), _ptr(wx), _ptr(wxx), _ptr(wt),
                                             _ptr(res), _ptr(div), _ptr(loss), _ptr(ws), _stream()),
              "pigs_pde_ns_residual_forward")
    return loss, res, div


def pde_ns_residual_backward(u, ux, wx, wxx, wt, nu, dt, g_loss, need=(True,) * 5):
    """-> (dL_du, dL_dux, dL_dwx, dL_dwxx, dL_dwt), None where not needed."""
    L = _lib.lib()
    N, u, ux, wx, wxx, wt = _ns_args(u, ux, wx, wxx, wt)
    g_loss = _f32c(g_loss, "dL_dloss").reshape(1)
    outs = [torch.empty_like(t) if n else None for t, n in zip((u, ux, wx, wxx, wt), need)]
    if N > 0:
        check(L.pigs_pde_ns_residual_backward(N, float(nu), float(dt), _ptr(u), _ptr(ux), _ptr(wx), _ptr(wxx), _ptr(wt),
                                              _ptr(g_loss), *[_ptr(o) for o in outs], _stream()),
              "pigs_pde_ns_residual_backward")
    return tuple(outs)


def _ts_args(ts, N):
    """ts: [N] tensor (TRAPEZOID) or a python number (0 = FORWARD rule, 1 = BACKWARD rule) -> (tensor or None, constant)."""
    if isinstance(ts, torch.Tensor):
        ts = _f32c(ts, "time_samples")
        if ts.numel() != N:
            raise RuntimeError("time_samples must hold one value per sample")
        return ts, 0.0
    return None, float(ts)


def _levels(name, cur, prev):
    cur, prev = _f32c(cur, name + "[-1]"), _f32c(prev, name + "[-2]")
    if cur.shape != prev.shape:
        raise RuntimeError("the two time levels of %s must have the same shape" % name)
    return cur, prev


def pde_residual_timelevels_forward(problem, ts, u, ux, lap, nu, dt, want_residual=False):
    """u, ux, lap: pairs (level [-1], level [-2]); ux may be None unless BURGERS.  -> (loss, residual or None)."""
    L = _lib.lib()
    u1, u0 = _levels("u", *u)
    lap1, lap0 = _levels("lap", *lap)
    ux1, ux0 = _levels("ux", *ux) if (ux is not None and problem == PDE_BURGERS) else (None, None)
    N, D, C, _, _, _, _ = _pde_args(problem, u1, ux1, lap1, u0)
    ts_t, ts_c = _ts_args(ts, N)
    dev = lap1.device
    loss = torch.zeros((), dtype=torch.float32, device=dev)
    res = torch.empty_like(lap1) if want_residual else None
    if N > 0:
        ws = _bytes(max(int(L.pigs_pde_residual_workspace_bytes(N, C)), 16), dev)
        check(L.pigs_pde_residual_timelevels_forward(problem, N, D, C, float(nu), float(dt), _ptr(ts_t), ts_c, _ptr(u1),
                                                     _ptr(u0), _ptr(ux1), _ptr(ux0), _ptr(lap1), _ptr(lap0), _ptr(res),
                                                     _ptr(loss), _ptr(ws), _stream()), "pigs_pde_residual_timelevels_forward")
    return loss, res


def pde_residual_timelevels_backward(problem, ts, u, ux, lap, nu, dt, g_loss, need=(True,) * 6):
    """-> (dL_du1, dL_du0, dL_dux1, dL_dux0, dL_dlap1, dL_dlap0); None where not needed (ux: BURGERS only)."""
    L = _lib.lib()
    u1, u0 = _levels("u", *u)
    lap1, lap0 = _levels("lap", *lap)
    ux1, ux0 = _levels("ux", *ux) if (ux is not None and problem == PDE_BURGERS) else (None, None)
    N, D, C, _, _, _, _ = _pde_args(problem, u1, ux1, lap1, u0)
    ts_t, ts_c = _ts_args(ts, N)
    g_loss = _f32c(g_loss, "dL_dloss").reshape(1)
    like = (u1, u0, ux1, ux0, lap1, lap0)
    outs = [torch.empty_like(t) if (n and t is not None) else None for t, n in zip(like, need)]
    if N > 0:
        check(L.pigs_pde_residual_timelevels_backward(problem, N, D, C, float(nu), float(dt), _ptr(ts_t), ts_c, _ptr(u1),
                                                      _ptr(u0), _ptr(ux1), _ptr(ux0), _ptr(lap1), _ptr(lap0), _ptr(g_loss),
                                                      *[_ptr(o) for o in outs], _stream()),
              "pigs_pde_residual_timelevels_backward")
    return tuple(outs)


def _ns_levels(u, ux, wx, wxx, w):
    cur = _ns_args(u[0], ux[0], wx[0], wxx[0], w[0])
    prev = _ns_args(u[1], ux[1], wx[1], wxx[1], w[1])
    if cur[0] != prev[0]:
        raise RuntimeError("the two time levels must hold the same number of samples")
    flat = []
    for a, b in zip(cur[1:], prev[1:]):
        flat += [a, b]
    return cur[0], flat  # N, [u1, u0, ux1, ux0, wx1, wx0, wxx1, wxx0, w1, w0]


def pde_ns_residual_timelevels_forward(ts, u, ux, wx, wxx, w, nu, dt, want_residual=False):
    """u, ux, wx, wxx, w: pairs (level [-1], level [-2]).  -> (loss, residual [N] or None, divergence [N] or None)."""
    L = _lib.lib()
    N, flat = _ns_levels(u, ux, wx, wxx, w)
    ts_t, ts_c = _ts_args(ts, N)
    dev = flat[0].device
    loss = torch.zeros((), dtype=torch.float32, device=dev)
    res = torch.empty(N, dtype=torch.float32, device=dev) if want_residual else None
    div = torch.empty(N, dtype=torch.float32, device=dev) if want_residual else None
    if N > 0:
        ws = _bytes(max(int(L.pigs_pde_residual_workspace_bytes(N, 2)), 16), dev)
        check(L.pigs_pde_ns_residual_timelevels_forward(N, float(nu), float(dt), _ptr(ts_t), ts_c, *[_ptr(t) for t in flat],
                                                        _ptr(res), _ptr(div), _ptr(loss), _ptr(ws), _stream()),
              "pigs_pde_ns_residual_timelevels_forward")
    return loss, res, div


def pde_ns_residual_timelevels_backward(ts, u, ux, wx, wxx, w, nu, dt, g_loss, need=(True,) * 10):
    """-> gradients in the order (u1, u0, ux1, ux0, wx1, wx0, wxx1, wxx0, w1, w0), None where not needed."""
    L = _lib.lib()
    N, flat = _ns_levels(u, ux, wx, wxx, w)
    ts_t, ts_c = _ts_args(ts, N)
    g_loss = _f32c(g_loss, "dL_dloss").reshape(1)
    outs = [torch.empty_like(t) if n else None for t, n in zip(flat, need)]
    if N > 0:
        check(L.pigs_pde_ns_residual_timelevels_backward(N, float(nu), float(dt), _ptr(ts_t), ts_c, *[_ptr(t) for t in flat],
                                                         _ptr(g_loss), *[_ptr(o) for o in outs], _stream()),
              "pigs_pde_ns_residual_timelevels_backward")
    return tuple(outs)


def vorticity_forward(x):
    """x [..., 2, 2] (trailing axes: spatial axis, channel; C = 2) -> x[..., 0, 1] - x[..., 1, 0], shape x.shape[:-2]."""
    L = _lib.lib()
    x = _al16(_f32c(x, "x"))
    if x.dim() < 2 or x.shape[-1] != 2 or x.shape[-2] != 2:
        raise RuntimeError("vorticity: trailing axes must be [2 (spatial), 2 (channels)]")
    out = torch.empty(x.shape[:-2], dtype=torch.float32, device=x.device)
    if out.numel() > 0:
        check(L.pigs_vorticity_forward(out.numel(), _ptr(x), _ptr(out), _stream()), "pigs_vorticity_forward")
    return out


def vorticity_backward(g):
    L = _lib.lib()
    g = _f32c(g, "dL_dout")
    d = torch.empty(tuple(g.shape) + (2, 2), dtype=torch.float32, device=g.device)
    if g.numel() > 0:
        check(L.pigs_vorticity_backward(g.numel(), _ptr(g), _ptr(d), _stream()), "pigs_vorticity_backward")
    return d


# ---- dense all-pairs path (gaussians.py:48-58, 89-101, 103-116) -------------------
def _dense_args(means, conics, values, samples):
    means = _f32c(means, "means")
    n, D = means.shape
    conics = _f32c(conics, "conics")
    values = _f32c(values, "values")
    samples = _f32c(samples, "samples")
    if D not in (1, 2) or conics.numel() != n * D * D:
        raise RuntimeError("dense path: means [n, D] with D in (1, 2) and FULL conics [n, D, D] expected")
    C = values.numel() // n if n > 0 else (values.shape[-1] if values.dim() > 1 else 1)
    if values.numel() != n * C or not 1 <= C <= 4:
        raise RuntimeError("dense path: values [n, C] with C <= 4 expected")
    N = samples.numel() // D
    return n, D, C, N, means, conics, values, samples


def dense_forward(fn_mask, means, conics, values, samples):
    """-> (u [N,C] | None, ux [N,D,C] | None, uxx [N,D,D,C] | None) for the bits of fn_mask (FN_VALUE | FN_DERIV | FN_HESS)."""
    L = _lib.lib()
    n, D, C, N, means, conics, values, samples = _dense_args(means, conics, values, samples)
    dev = means.device
    mk = lambda bit, *sh: torch.zeros(sh, dtype=torch.float32, device=dev) if fn_mask & bit else None
    u, ux, uxx = mk(1, N, C), mk(2, N, D, C), mk(4, N, D, D, C)
    if n > 0 and N > 0:
        check(L.pigs_dense_forward(fn_mask, n, D, C, N, _ptr(means), _ptr(conics), _ptr(values), _ptr(samples), _ptr(u),
                                   _ptr(ux), _ptr(uxx), _stream()), "pigs_dense_forward")
    return u, ux, uxx


def dense_backward(fn_mask, means, conics, values, samples, g_u, g_ux, g_uxx):
    """-> (dL_dmeans [n,D], dL_dconics [n,D,D], dL_dvalues [n,C])."""
    L = _lib.lib()
    n, D, C, N, means, conics, values, samples = _dense_args(means, conics, values, samples)
    gs = [None if (g is None or not fn_mask & bit) else _f32c(g, "dL_dout") for g, bit in ((g_u, 1), (g_ux, 2), (g_uxx, 4))]
    dm, dc, dv = torch.zeros_like(means), torch.zeros((n, D, D), dtype=torch.float32, device=means.device), \
        torch.zeros((n, C), dtype=torch.float32, device=means.device)
    if n > 0 and N > 0:
        check(L.pigs_dense_backward(fn_mask, n, D, C, N, _ptr(means), _ptr(conics), _ptr(values), _ptr(samples),
                                    *[_ptr(g) for g in gs], _ptr(dm), _ptr(dc), _ptr(dv), _stream()), "pigs_dense_backward")
    return dm, dc, dv


# ---- compiled binding ------------------------------------------------------------------------------------------------
# pigs_b200/_Cext.so (csrc/torch_ext.cpp) implements the entry points above in C++ (the reference's own binding is a
# pybind module too, ext.cpp:19-32).  When it is built it takes over the hot entry points; the ctypes functions above stay
# importable as `<name>_ctypes` (torch-free-build alternative, and the cross-check in tests/test_compiled_binding.py).
# PIGS_B200_BINDING=ctypes keeps the ctypes path.
_HOT = ("preprocess_gaussians", "sample_forward", "sample_backward", "preprocess_aggregate", "aggregate_neighbors",
        "aggregate_neighbors_backward", "sample_gaussians", "sample_gaussians_backward", "sample_gaussians_derivative",
        "sample_gaussians_derivative_backward", "sample_gaussians_laplacian", "sample_gaussians_laplacian_backward",
        "sample_gaussians_third_derivative", "sample_gaussians_third_derivative_backward")
BINDING = "ctypes"
_ext = None


def _load_compiled():
    import importlib.util
    import os
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_Cext.so")
    if os.environ.get("PIGS_B200_BINDING", "compiled") == "ctypes" or not os.path.exists(path):
        return None
    _lib.lib()  # the one libpigs_b200.so instance both bindings share
    spec = importlib.util.spec_from_file_location("pigs_b200._Cext", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def use_binding(name):
    """Select "compiled" or "ctypes" for the hot entry points (tests compare the two)."""
    global BINDING, _ext
    g = globals()
    if name == "compiled":
        if _ext is None:
            _ext = _load_compiled()
        if _ext is None:
            raise RuntimeError("pigs_b200/_Cext.so is not built (python -c 'import __graft_entry__ as g; g.build()')")
        for n in _HOT:
            g[n] = _wrap_compiled(n)
    elif name == "ctypes":
        for n in _HOT:
            g[n] = g[n + "_ctypes"]
    else:
        raise ValueError(name)
    BINDING = name


def _wrap_compiled(n):
    f = getattr(_ext, n)
    if n == "sample_forward":
        return lambda *a: tuple(f(*a))
    if n == "sample_backward":
        def bw(fn_mask, means, values, conics, samples, num_rendered, grads, *rest):
            return f(fn_mask, means, values, conics, samples, num_rendered, list(grads), *rest)
        return bw
    if n == "preprocess_gaussians":
        def pre(means, values, covariances, conics, samples, debug, bounds=None, cull=None):
            if cull is not None:
                cull = (float(cull[0]), int(cull[1]))
            return f(means, values, covariances, conics, samples, bool(debug), bounds, cull)
        return pre
    if n in ("aggregate_neighbors", "aggregate_neighbors_backward"):
        slow = globals()[n + "_ctypes"]

        def agg(features, *rest):
            if features.dtype == torch.float64:  # the double instantiation is bound through ctypes only
                return slow(features, *rest)
            return f(features, *rest)
        return agg
    if n == "preprocess_aggregate":
        def pa(means, conics, radii, debug, rows=None):
            if rows is not None:
                rows = (int(rows[0]), int(rows[1]))
            return f(means, conics, radii, bool(debug), rows)
        return pa
    return f


for _n in _HOT:
    globals()[_n + "_ctypes"] = globals()[_n]
try:
    _ext = _load_compiled()
    if _ext is not None:
        use_binding("compiled")
except Exception as _ex:  # a stale or unloadable _Cext.so must not take the package down: the ctypes binding is complete
    import warnings
    warnings.warn("pigs_b200: compiled binding not usable (%s); using the ctypes binding" % (_ex,))
    _ext = None
