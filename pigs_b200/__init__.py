"""pigs_b200 — B200-native (sm_100a) implementation of kr4b/pigs' differentiable Gaussian-mixture
sampling path, behind the reference's own Python API.

Public surface = diff-gaussian-sampling/diff_gaussian_sampling/__init__.py (reference), name for name:
  GaussianSampler(debug) with .preprocess / .sample_gaussians / .sample_gaussians_derivative /
  .sample_gaussians_laplacian / .sample_gaussians_third_derivative / .preprocess_aggregate /
  .aggregate_neighbors, and the module-level functions sample_gaussians(...), ...,
  preprocess_gaussians, preprocess_aggregate, aggregate_neighbors.
Extensions (additive, keyword-only): GaussianSampler(debug, fused=...), .sample_fused(mask), .pair_count(),
.sample_laplacian() (trace of the Hessian, [N, C] — what model_pn.py:613-625 pde_rhs consumes; pigs_b200/pde.py).

Everything below the API is hand-written CUDA in libpigs_b200.so (include/pigs_b200.h); there is no
CPU or PyTorch fallback.
"""
import os

import torch

from . import _C
from ._C import FN_DERIV, FN_HESS, FN_LAPL, FN_THIRD, FN_VALUE

__all__ = [
    "GaussianSampler", "sample_gaussians", "sample_gaussians_derivative", "sample_gaussians_laplacian",
    "sample_gaussians_third_derivative", "aggregate_neighbors", "preprocess_gaussians", "preprocess_aggregate",
    "FN_VALUE", "FN_DERIV", "FN_HESS", "FN_THIRD", "FN_LAPL", "set_deterministic", "get_deterministic",
]


def set_deterministic(on):
    """(extension) Deterministic forward: bitwise run-to-run reproducible u / grad u / Hessian, like the reference's
    forward (include/pigs_b200.h pigs_set_deterministic).  Takes effect at the next preprocess().  Also switched on by
    the environment variable PIGS_DETERMINISTIC=1 when the library is loaded."""
    from . import _lib
    _lib.lib().pigs_set_deterministic(1 if on else 0)


def get_deterministic():
    from . import _lib
    return bool(_lib.lib().pigs_get_deterministic())


def cpu_deep_copy_tuple(input_tuple):
    return tuple(item.cpu().clone() if isinstance(item, torch.Tensor) else item for item in input_tuple)


def call_debug(func, debug, name, *args):
    """Reference debug semantics (__init__.py:38-50): with debug on, inputs are copied to the CPU before the
    call and dumped to snapshot_<name>.dump if it raises; the native side synchronises after each launch."""
    if debug:
        cpu_args = cpu_deep_copy_tuple(args)
        try:
            results = func(*args)
        except Exception as ex:
            torch.save(cpu_args, "snapshot_{}.dump".format(name))
            print("\nAn error occured in {}. Please forward snapshot_{}.dump for debugging.".format(name, name))
            raise ex
    else:
        results = func(*args)
    return results


def preprocess_gaussians(means, values, covariances, conics, samples, debug):
    return call_debug(_C.preprocess_gaussians, debug, "preprocess", means, values, covariances, conics, samples, debug)


def preprocess_aggregate(means, conics, radii, debug):
    return call_debug(_C.preprocess_aggregate, debug, "preprocess_agg", means, conics, radii, debug)


class _SampleFused(torch.autograd.Function):
    """One autograd node for any subset of {value, derivative, Hessian, third}: one fused forward kernel,
    one fused backward kernel.  Inputs are kept on ctx (not save_for_backward) so that, like the
    reference's separate nodes, each output can be back-propagated in its own backward() call."""

    @staticmethod
    def forward(ctx, mask, means, values, conics, samples, num_rendered, binning_buffer, sample_binning_buffer, ranges,
                sample_ranges, debug, reduce=None):
        args = (means, values, conics, samples, num_rendered, binning_buffer, sample_binning_buffer, ranges,
                sample_ranges, debug)
        outs = call_debug(lambda *a: _C.sample_forward(mask, *a), debug, "fw%d" % mask, *args)
        ctx.mask = mask
        ctx.args = args
        ctx.versions = _versions(args)
        ctx.reduce = reduce  # sharded sampler: callable applied to the packed gradient buffer (NCCL all-reduce)
        # outputs nobody back-propagates through arrive as None instead of materialised zeros and drop out of the backward
        ctx.set_materialize_grads(False)
        return tuple(o for o in outs if o is not None)

    @staticmethod
    def backward(ctx, *grad_outs):
        mask = ctx.mask
        _check_versions(ctx.args, ctx.versions)
        means, values, conics, samples, num_rendered, bb, sbb, ranges, sample_ranges, debug = ctx.args
        grads, it = [None] * 4, iter(grad_outs)
        live = 0
        for k in range(4):
            bit = _slot_bit(mask, k)
            if bit:
                g = next(it)
                if g is not None:
                    grads[k] = g
                    live |= bit
        if live == 0:
            return (None,) * 12
        dm, dv, dc = call_debug(lambda *a: _C.sample_backward(live, *a), debug, "bw%d" % mask, means, values, conics,
                                samples, num_rendered, grads, bb, sbb, ranges, sample_ranges, debug)
        if ctx.reduce is not None:
            ctx.reduce(_C.packed_base(dm))  # dm, dv, dc are views of one flat buffer: one collective
        return (None, dm, dv, dc, None, None, None, None, None, None, None, None)


def _versions(args):
    return tuple(a._version if isinstance(a, torch.Tensor) else None for a in args)


def _check_versions(args, versions):
    """The inputs are kept on ctx (not save_for_backward) so that each output can be back-propagated in its own
    backward() call like the reference's separate nodes; autograd's in-place check is re-created here."""
    for a, v in zip(args, versions):
        if isinstance(a, torch.Tensor) and a._version != v:
            raise RuntimeError("one of the variables needed for gradient computation has been modified by an inplace "
                               "operation (pigs_b200: an input of the sampling / aggregation call changed between its "
                               "forward and its backward)")


def _slot_bit(mask, k):
    """The function bit of `mask` whose output sits in slot k of the (value, derivative, Hessian | Laplacian, third)
    4-tuple, or 0."""
    if k == 2 and (mask & FN_LAPL):
        return FN_LAPL
    return mask & (1 << k)


def _sample_mask(mask, *args):
    return _SampleFused.apply(mask, *args)


# module-level functions, reference signatures (__init__.py:21-36): the first positional parameter is
# called `debug` there but receives `means`; args = (means, values, conics, samples, num_rendered,
# binning_buffer, sample_binning_buffer, ranges, sample_ranges, debug)
def sample_gaussians(debug, *args):
    return _sample_mask(FN_VALUE, debug, *args)[0]


def sample_gaussians_derivative(debug, *args):
    return _sample_mask(FN_DERIV, debug, *args)[0]


def sample_gaussians_laplacian(debug, *args):
    return _sample_mask(FN_HESS, debug, *args)[0]


def sample_gaussians_third_derivative(debug, *args):
    return _sample_mask(FN_THIRD, debug, *args)[0]


class _AggregateNeighbors(torch.autograd.Function):
    @staticmethod
    def forward(ctx, features, transform, queries, keys, frequencies, distance_transform, indices, ranges, dists,
                densities, inv_total_densities, debug, reduce=None):
        args = (features, transform, queries, keys, frequencies, distance_transform, indices, ranges, dists, densities,
                inv_total_densities, debug)
        weights, embeddings, factors, neighbor_features = call_debug(_C.aggregate_neighbors, debug, "aggregate", *args)
        ctx.args = args
        ctx.versions = _versions(args)
        ctx.saved = (weights, embeddings, factors)
        ctx.reduce = reduce  # sharded sampler: sums the row-sharded output and the six gradients over ranks
        if reduce is not None:
            neighbor_features = reduce.output(neighbor_features)
        return neighbor_features

    @staticmethod
    def backward(ctx, grad_out):
        _check_versions(ctx.args, ctx.versions)
        (features, transform, queries, keys, frequencies, distance_transform, indices, ranges, dists, densities,
         inv_total_densities, debug) = ctx.args
        weights, embeddings, factors = ctx.saved
        grads = call_debug(_C.aggregate_neighbors_backward, debug, "aggregate_bw", features, transform, queries, keys,
                           frequencies, distance_transform, indices, ranges, dists, densities, weights, embeddings,
                           factors, inv_total_densities, grad_out.contiguous(), debug)
        if ctx.reduce is not None:
            grads = ctx.reduce.gradients(grads)
        return tuple(grads) + (None,) * 7


def aggregate_neighbors(features, transform, queries, keys, frequencies, distance_transform, indices, ranges, dists,
                        densities, inv_total_densities, debug, reduce=None):
    return _AggregateNeighbors.apply(features, transform, queries, keys, frequencies, distance_transform, indices,
                                     ranges, dists, densities, inv_total_densities, debug, reduce)


_BIT_OF = {"u": FN_VALUE, "du": FN_DERIV, "ddu": FN_HESS, "dddu": FN_THIRD}


class GaussianSampler:
    """Stateful sampler, reference API (__init__.py:214-317).

    fused (extension): True / False / "auto" (default, env PIGS_B200_FUSE=0 disables).  The reference's
    callers ask for u, grad u and the Hessian in three calls after one preprocess(); with fusion on, the
    sampler remembers which outputs were requested during the last preprocess() cycle WITH THE SAME NUMBER OF SAMPLES
    (the call site: interior / boundary / means cycles alternate on one sampler in model_pn.py) and, on the first
    request of the next such cycle, evaluates that whole set in ONE kernel (and back-propagates it in one), handing
    out the cached pieces.  Results are identical to the unfused calls."""

    def __init__(self, debug, fused="auto", cull=None):
        self.debug = debug
        # (extension) opt-in culled mode: cull = sigmas or (sigmas, subdiv).  A SUBSET of the
        # reference's pair set: every reference tile is cut subdiv x subdiv and pairs whose fine tile is farther than
        # `sigmas` standard deviations (major axis) from the mean are dropped, each worth < exp(-sigmas^2/2) of its
        # value.  Default None = the reference's pair set, the mode every parity claim is made in.
        self.cull = cull
        if fused == "auto":
            fused = os.environ.get("PIGS_B200_FUSE", "1") != "0"
        self.fused = bool(fused)
        # The plan is keyed by the call site, identified by the number of samples: the reference driver alternates
        # cycles on ONE sampler (model_pn.py:768-786: interior samples -> u, ux, uxx; boundary samples -> u only;
        # Model.forward: samples = means -> u, ux, uxx), and each kind of cycle keeps asking for its own set.
        self._plans = {}    # samples per cycle -> mask requested during the last cycle with that many samples
        self._key = None
        self._plan = 0      # mask the current cycle is expected to ask for
        self._seen = 0      # mask requested so far in this cycle
        self._cache = {}    # bit -> tensor produced by a fused call in this cycle
        self._cache_grad = None  # torch.is_grad_enabled() when the cache was filled

    def preprocess(self, means, values, covariances, conics, samples, bounds=None):
        debug = self.debug
        cull = self._cull_params(covariances, means.shape[-1]) if self.cull is not None else None
        if bounds is None and cull is None:
            res = preprocess_gaussians(means, values, covariances, conics, samples, debug)
        else:
            res = call_debug(lambda *a: _C.preprocess_gaussians(*a, bounds=bounds, cull=cull), debug, "preprocess", means,
                             values, covariances, conics, samples, debug)
        num_rendered, binning_buffer, sample_binning_buffer, ranges, sample_ranges, radii = res
        self.means = means
        self.values = values
        self.conics = conics
        self.samples = samples
        self.num_rendered = num_rendered
        self.binning_buffer = binning_buffer
        self.sample_binning_buffer = sample_binning_buffer
        self.ranges = ranges
        self.sample_ranges = sample_ranges
        self.radii = radii
        self._sample_state = (num_rendered, binning_buffer, sample_binning_buffer, ranges, sample_ranges)
        if self._key is not None and self._seen:
            self._plans[self._key] = self._seen
        self._key = int(samples.shape[0])
        self._plan, self._seen, self._cache = self._plans.get(self._key, 0), 0, {}
        if (self._plan & FN_HESS) and (self._plan & FN_LAPL):
            self._plan &= ~FN_LAPL  # shared output slot: the Laplacian request then runs on its own

    def _cull_params(self, covariances, D):
        """(sigmas, subdiv).  Without an explicit subdivision: fine tiles about half as wide as the mean cut radius
        (1 <= subdiv <= 16), which keeps the overshoot of the box test moderate without exploding the instance count.
        Costs one extra device->host scalar read per preprocess."""
        if isinstance(self.cull, (tuple, list)):
            return float(self.cull[0]), int(self.cull[1])
        sig = float(self.cull)
        K = D * (D + 1) // 2
        cov = covariances.detach().reshape(-1, K)
        var = cov[:, 0] if D == 1 else torch.maximum(cov[:, 0], cov[:, 2])
        rc = sig * float(var.sqrt().mean().item()) if cov.shape[0] else 1.0
        return sig, int(min(16, max(1, round(2.0 * 0.51 / max(rc, 1e-6)))))

    def _args(self):
        nr, bb, sbb, rg, srg = self._sample_state
        reduce = getattr(self, "_reduce", None)
        if reduce is not None:
            group = self.group
            return (self.means, self.values, self.conics, self.samples, nr, bb, sbb, rg, srg, self.debug,
                    lambda flat: reduce(flat, group))
        return (self.means, self.values, self.conics, self.samples, nr, bb, sbb, rg, srg, self.debug)

    def sample_fused(self, mask):
        """(extension) evaluate every function in `mask` with one fused kernel; returns a tuple in
        (value, derivative, Hessian, third) order restricted to the requested ones."""
        return _sample_mask(mask, *self._args())

    def _request(self, bit):
        self._seen |= bit
        if bit in self._cache:
            out = self._cache.pop(bit)
            if self._cache_grad == torch.is_grad_enabled():
                return out
            return self.sample_fused(bit)[0]  # fused under another grad mode: its pieces carry the wrong (or no) graph
        if self.fused and (self._plan & bit) and self._plan != bit and not self._cache:
            self._cache_grad = torch.is_grad_enabled()
            outs = iter(self.sample_fused(self._plan))
            for k in range(4):
                b = _slot_bit(self._plan, k)
                if b:
                    self._cache[b] = next(outs)
            return self._cache.pop(bit)
        return self.sample_fused(bit)[0]

    def sample_gaussians(self):
        return self._request(FN_VALUE)

    def sample_gaussians_derivative(self):
        return self._request(FN_DERIV)

    def sample_gaussians_laplacian(self):
        return self._request(FN_HESS)

    def sample_gaussians_third_derivative(self):
        return self._request(FN_THIRD)

    def sample_laplacian(self):
        """(extension) Laplacian of the mixture at every sample, [N, C]: the trace
        `uxx[:,0,0] + uxx[:,1,1]` that pde_rhs takes of sample_gaussians_laplacian()'s [N, D, D, C] Hessian
        (model_pn.py:613-625), formed in the evaluation kernel's epilogue so neither the Hessian nor its upstream
        gradient goes through HBM.  Not combinable with sample_gaussians_laplacian() in one preprocess cycle's
        fused plan (the two share an output slot); asking for both simply runs them unfused."""
        return self._request(FN_LAPL)

    def pair_weights(self):
        """(extension) per-sample size of the reference pair set: the number of Gaussians binned to the sample's tile,
        int64 [N] (sums to pair_count()).  Used to cut sample shards of equal WORK (sharded.balanced_ranges)."""
        ranges, sample_ranges = self._sample_state[3], self._sample_state[4]
        N = self.samples.shape[0]
        b = ranges.numel() // 8 - 1
        w = torch.zeros(N, dtype=torch.int64, device=ranges.device)
        if b <= 0 or N == 0:
            return w
        g = ranges.view(torch.int32)[: 2 * b].view(b, 2).to(torch.int64)
        sr = sample_ranges.view(torch.int32)[: 2 * b].view(b, 2).to(torch.int64)
        _, _, sl = _C.point_lists(self.num_rendered, self.binning_buffer, self.sample_binning_buffer, N)
        # sample_point_list is tile-major: tile t owns positions [sr[t,0], sr[t,1])
        ns = sr[:, 1] - sr[:, 0]
        first = torch.cumsum(ns, 0) - ns                       # index of each tile's first entry in the concatenation
        pos = torch.repeat_interleave(sr[:, 0] - first, ns) + torch.arange(int(ns.sum()), device=ns.device)
        per_pos = torch.zeros(N, dtype=torch.int64, device=ranges.device)
        per_pos[pos] = torch.repeat_interleave(g[:, 1] - g[:, 0], ns)   # Gaussians of the tile, once per sample of the tile
        w[sl.to(torch.int64)] = per_pos
        return w

    def pair_count(self):
        """(extension) size of the reference pair set: sum over tiles of n_samples(tile) * n_gaussians(tile)."""
        ranges, sample_ranges = self._sample_state[3], self._sample_state[4]
        b = ranges.numel() // 8 - 1
        if b <= 0:
            return 0
        g = ranges.view(torch.int32)[: 2 * b].view(b, 2).to(torch.int64)
        s = sample_ranges.view(torch.int32)[: 2 * b].view(b, 2).to(torch.int64)
        return int(((g[:, 1] - g[:, 0]) * (s[:, 1] - s[:, 0])).sum().item())

    def preprocess_aggregate(self):
        indices, ranges, dists, densities, inv_total_densities = preprocess_aggregate(self.means, self.conics, self.radii,
                                                                                       self.debug)
        self.indices = indices
        self.ranges = ranges  # NB: like the reference (__init__.py:298) this overwrites the tile ranges attribute
        self.dists = dists
        self.densities = densities
        self.inv_total_densities = inv_total_densities

    def aggregate_neighbors(self, features, transform, queries, keys, frequencies, distance_transform):
        return aggregate_neighbors(features, transform, queries, keys, frequencies, distance_transform, self.indices,
                                   self.ranges, self.dists, self.densities, self.inv_total_densities, self.debug)
