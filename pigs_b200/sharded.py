"""Multi-GPU data parallelism over collocation points: one process per GPU, the samples are
split into contiguous index ranges, the Gaussians (and their binning) are replicated, and the parameter
gradients are summed with ONE NCCL all-reduce of a packed [P*(D+C+D(D+1)/2)] fp32 buffer per backward.
The tile grid must be the single-GPU one on every rank, so the sample bounding box (the reference computes it
from the samples on every call, sample_points.cu:70-71) is all-reduced first: 2*D floats, one collective.

The reference has no distributed code of any kind; this is additive.  The three helpers below are pure
torch.distributed host logic (tested on CPU with gloo, world_size 2); the compute stays in libpigs_b200.so.
"""
import torch
import torch.distributed as dist

from . import GaussianSampler, _C, aggregate_neighbors


def shard_range(N, rank, world):
    """Contiguous even split of N samples: ranks [0, N % world) get one extra."""
    base, rem = divmod(int(N), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def balanced_ranges(weights, world):
    """Contiguous split of len(weights) rows into `world` windows of (nearly) equal total weight -> [(lo, hi)] * world.
    With weights = GaussianSampler.pair_weights() of a pass over ALL samples this equalises the evaluation work per rank
    where an even split does not (model_pn.py lays its Gaussians out as a grid, so a contiguous range of the means is a
    band of the domain and the bands differ in pairs per sample).  Deterministic, so every rank computes the same cuts."""
    w = weights.reshape(-1).to(torch.float64)
    n, world = w.numel(), int(world)
    if n == 0:
        return [(0, 0)] * world
    cum = torch.cumsum(w, 0)
    total = float(cum[-1])
    if total <= 0.0:
        return [shard_range(n, r, world) for r in range(world)]
    targets = torch.arange(1, world, dtype=torch.float64, device=w.device) * (total / world)
    cuts = torch.searchsorted(cum, targets, right=False).clamp_(0, n).tolist() if world > 1 else []
    edges = [0] + [int(c) for c in cuts] + [n]
    for i in range(1, len(edges)):  # monotone
        edges[i] = max(edges[i], edges[i - 1])
    return [(edges[r], edges[r + 1]) for r in range(world)]


def allreduce_bounds(mm, D, group=None):
    """mm = float32[2*D] {min[0..D), max[0..D)} of the local samples -> global bounds, in place, one collective
    (MAX over {-min, max})."""
    if group is None and not dist.is_initialized():
        return mm
    mm[:D].neg_()
    dist.all_reduce(mm, op=dist.ReduceOp.MAX, group=group)
    mm[:D].neg_()
    return mm


def allreduce_grads(flat, group=None):
    """Sum of the packed parameter-gradient buffer over ranks, in place."""
    if group is None and not dist.is_initialized():
        return flat
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    return flat


def pack_grads(grads):
    """Six aggregation gradients (any may be None) -> one flat fp32 buffer + the recipe to split it again."""
    live = [(k, g) for k, g in enumerate(grads) if g is not None]
    flat = torch.cat([g.reshape(-1) for _, g in live]) if live else None
    return flat, [(k, g.shape) for k, g in live]


def unpack_grads(flat, recipe, n):
    out, at = [None] * n, 0
    for k, shape in recipe:
        cnt = 1
        for d in shape:
            cnt *= int(d)
        out[k] = flat[at:at + cnt].view(shape)
        at += cnt
    return out


class AggregateReduce:
    """Collectives of the row-sharded aggregation: each rank owns the neighbour lists of a
    contiguous range of query Gaussians, so its forward output has the other ranks' rows zero (one all-reduce
    assembles [P, L]) and its backward holds the contributions of its own rows to all six gradients (one all-reduce
    of the packed buffer)."""

    def __init__(self, group=None):
        self.group = group

    def output(self, neighbor_features):
        return allreduce_grads(neighbor_features, self.group)

    def gradients(self, grads):
        flat, recipe = pack_grads(grads)
        if flat is None:
            return grads
        allreduce_grads(flat, self.group)
        return unpack_grads(flat, recipe, len(grads))


class ShardedGaussianSampler(GaussianSampler):
    """GaussianSampler whose `samples` are this rank's shard.  Same API; preprocess() bins on the global tile
    grid, backward all-reduces the gradients w.r.t. means / values / conics (so every rank ends up with the
    full-batch gradient, as a single-GPU run over all samples would)."""

    def __init__(self, debug, group=None, fused="auto", cull=None):
        super().__init__(debug, fused=fused, cull=cull)
        self.group = group
        self._reduce = allreduce_grads

    def preprocess(self, means, values, covariances, conics, samples):
        D = means.size(-1)
        mm = _C.sample_bounds(samples, D)
        allreduce_bounds(mm, D, self.group)
        super().preprocess(means, values, covariances, conics, samples, bounds=mm)

    def _rank_world(self):
        if not dist.is_initialized():
            return 0, 1
        return dist.get_rank(self.group), dist.get_world_size(self.group)

    def preprocess_aggregate(self, row_range=None):
        """Neighbour lists of this rank's contiguous range of query Gaussians (every Gaussian stays a candidate
        neighbour); the other rows are empty.  1/world of the CSR per GPU.  row_range = (lo, hi) overrides the even
        split, e.g. with this rank's window of balanced_ranges()."""
        rank, world = self._rank_world()
        self.row_range = shard_range(self.means.size(0), rank, world) if row_range is None else tuple(row_range)
        (self.indices, self.ranges, self.dists, self.densities,
         self.inv_total_densities) = _C.preprocess_aggregate(self.means, self.conics, self.radii, self.debug,
                                                             rows=self.row_range)

    def aggregate_neighbors(self, features, transform, queries, keys, frequencies, distance_transform):
        """Full [P, L] output on every rank; gradients of all six inputs summed over ranks."""
        return aggregate_neighbors(features, transform, queries, keys, frequencies, distance_transform, self.indices,
                                   self.ranges, self.dists, self.densities, self.inv_total_densities, self.debug,
                                   reduce=AggregateReduce(self.group))
