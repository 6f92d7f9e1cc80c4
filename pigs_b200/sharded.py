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

from . import GaussianSampler, _C


def shard_range(N, rank, world):
    """Contiguous even split of N samples: ranks [0, N % world) get one extra."""
    base, rem = divmod(int(N), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


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
