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


def _rank_world(group=None):
    if not dist.is_initialized():
        return 0, 1
    return dist.get_rank(group), dist.get_world_size(group)


def _global_rank(group, r):
    return r if group is None else dist.get_global_rank(group, r)


def gather_windows(window, device, group=None):
    """Every rank's (lo, hi) row window -> list of tuples, identical on all ranks (one small all_gather)."""
    rank, world = _rank_world(group)
    if world == 1:
        return [tuple(int(v) for v in window)]
    mine = torch.tensor([int(window[0]), int(window[1])], dtype=torch.int64, device=device)
    parts = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(parts, mine, group=group)
    return [tuple(int(v) for v in t.tolist()) for t in parts]


def allgather_rows(buf, windows, group=None):
    """buf [P, W]: this rank has written its window's rows; on return every window holds its owner's rows.
    Equal contiguous windows -> ONE all-gather straight into `buf` (NCCL gathers in place); otherwise one broadcast
    per window.  (The all-reduce this replaces moved P rows of mostly zeros per rank.)"""
    rank, world = _rank_world(group)
    if world == 1:
        return buf
    sizes = {hi - lo for lo, hi in windows}
    tiled = windows[0][0] == 0 and windows[-1][1] == buf.size(0) and all(a[1] == b[0] for a, b in zip(windows, windows[1:]))
    if len(sizes) == 1 and tiled and buf.is_contiguous():
        lo, hi = windows[rank]
        mine = buf[lo:hi]
        if dist.get_backend(group) != "nccl":
            mine = mine.clone()  # only NCCL documents the in-place form
        dist.all_gather_into_tensor(buf, mine, group=group)
        return buf
    for r, (lo, hi) in enumerate(windows):
        if hi > lo:
            dist.broadcast(buf[lo:hi], src=_global_rank(group, r), group=group)
    return buf


class AggregateReduce:
    """Collectives of the row-sharded aggregation with REPLICATED results: each rank owns the
    neighbour lists of a contiguous window of query Gaussians and every rank ends up with the full [P, L] output and the
    full gradients.  Rows owned by exactly one rank (the output, dL/dqueries) are ALL-GATHERED window by window; only the
    gradients that really are sums over ranks (dL/dfeatures, dL/dkeys — scattered to neighbour rows — and the three small
    shared parameters) are all-reduced, packed into one buffer.  `windows` = every rank's (lo, hi); without it (legacy)
    everything is all-reduced."""

    ROW_OWNED = (2,)  # position of dL/dqueries in the six gradients

    def __init__(self, group=None, windows=None):
        self.group = group
        self.windows = windows

    def output(self, neighbor_features):
        if self.windows is None:
            return allreduce_grads(neighbor_features, self.group)
        return allgather_rows(neighbor_features, self.windows, self.group)

    def gradients(self, grads):
        grads = list(grads)
        owned = {}
        if self.windows is not None:
            for k in self.ROW_OWNED:
                if grads[k] is not None:
                    owned[k] = allgather_rows(grads[k], self.windows, self.group)
                    grads[k] = None
        flat, recipe = pack_grads(grads)
        if flat is not None:
            allreduce_grads(flat, self.group)
            grads = unpack_grads(flat, recipe, len(grads))
        for k, g in owned.items():
            grads[k] = g
        return grads


class RowHaloPlan:
    """Exchange plan of the row-PARALLEL aggregation (each rank keeps only its own rows of features / queries / keys,
    as a row-parallel MLP in front of the operator produces them, and gets back its own rows of the output and of the
    gradients).  The neighbour-indexed operands are needed for the rows j the window's CSR names: the strip
    [jlo, jhi) = window + halo.  With spatially local neighbour lists (model_pn.py lays the Gaussians out as a grid) the
    halo is a few thousand rows, so the exchange is a handful of small point-to-point copies instead of an all-gather /
    reduce-scatter of [P, L]; lists whose strips cover more than `dense_fraction` of P fall back to those collectives.
    Built once per preprocess_aggregate (one tiny all_gather + one host read of the index range); identical decisions on
    every rank because they are taken from the gathered table."""

    def __init__(self, P, window, indices, group=None, dense_fraction=0.5):
        self.P, self.group = int(P), group
        self.lo, self.hi = int(window[0]), int(window[1])
        self.rank, self.world = _rank_world(group)
        # index range of the window's lists (entries < 0 are the search's "skipped" slots), gathered from every rank
        # together with the windows: one collective and ONE device->host read per plan
        dev = indices.device
        if indices.numel():
            imin = torch.where(indices >= 0, indices, torch.full_like(indices, self.P)).amin().reshape(1)
            imax = (indices.amax() + 1).reshape(1)
        else:
            imin = torch.full((1,), self.P, dtype=torch.int64, device=dev)
            imax = torch.zeros(1, dtype=torch.int64, device=dev)
        mine = torch.cat([torch.tensor([self.lo, self.hi], dtype=torch.int64, device=dev), imin.to(torch.int64),
                          imax.to(torch.int64)])
        if self.world > 1:
            parts = [torch.zeros_like(mine) for _ in range(self.world)]
            dist.all_gather(parts, mine, group=group)
            rows = torch.stack(parts).cpu().tolist()
        else:
            rows = [mine.cpu().tolist()]
        # the strip always contains the window; a window without valid slots needs nothing else
        self.table = [(lo, hi, min(a, lo), max(b, hi)) if a < b else (lo, hi, lo, hi) for lo, hi, a, b in rows]
        self.jlo, self.jhi = self.table[self.rank][2], self.table[self.rank][3]
        self.windows = [(t[0], t[1]) for t in self.table]
        halo = []
        for r, (lo_r, hi_r, jlo_r, jhi_r) in enumerate(self.table):
            halo.append(sum(max(0, min(jhi_r, hi_s) - max(jlo_r, lo_s)) for s, (lo_s, hi_s, _, _) in enumerate(self.table) if s != r))
        self.dense = self.world > 1 and max(halo) > dense_fraction * self.P
        self.recv, self.send = [], []  # (peer, a, b): rows of the peer's window my strip needs / rows of mine its strip needs
        for s, (lo_s, hi_s, jlo_s, jhi_s) in enumerate(self.table):
            if s == self.rank:
                continue
            a, b = max(self.jlo, lo_s), min(self.jhi, hi_s)
            if a < b:
                self.recv.append((s, a, b))
            a, b = max(jlo_s, self.lo), min(jhi_s, self.hi)
            if a < b:
                self.send.append((s, a, b))
        self.halo_rows = halo[self.rank]

    def _exchange(self, ops):
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()

    def gather(self, bufs):
        """Forward: bufs = full-height [P, W] buffers holding this rank's window rows; fills the halo rows of the strip
        from their owners (dense fallback: every window)."""
        if self.world == 1:
            return bufs
        if self.dense:
            for b in bufs:
                allgather_rows(b, self.windows, self.group)
            return bufs
        ops = []
        for s, a, b in self.send:
            ops += [dist.P2POp(dist.isend, buf[a:b], _global_rank(self.group, s), self.group) for buf in bufs]
        for s, a, b in self.recv:
            ops += [dist.P2POp(dist.irecv, buf[a:b], _global_rank(self.group, s), self.group) for buf in bufs]
        self._exchange(ops)
        return bufs

    def scatter_add(self, bufs):
        """Backward: bufs = full-height [P, W] partial gradients this rank accumulated over its strip; on return each
        buffer's WINDOW rows hold the sum over ranks (the other rows are unspecified)."""
        if self.world == 1:
            return bufs
        if self.dense:
            for b in bufs:
                dist.all_reduce(b, group=self.group)
            return bufs
        ops, tmp = [], []
        for s, a, b in self.recv:  # rows of s's window that I touched go to s
            ops += [dist.P2POp(dist.isend, buf[a:b], _global_rank(self.group, s), self.group) for buf in bufs]
        for s, a, b in self.send:  # s touched rows [a, b) of my window
            for buf in bufs:
                t = torch.empty_like(buf[a:b])
                tmp.append((buf, a, b, t))
                ops.append(dist.P2POp(dist.irecv, t, _global_rank(self.group, s), self.group))
        self._exchange(ops)
        for buf, a, b, t in tmp:
            buf[a:b] += t
        return bufs


class _AggregateRows(torch.autograd.Function):
    """Row-parallel aggregation: own rows in, own rows out, halo exchange either side of the kernels, one small
    all-reduce for the three shared parameters' gradients."""

    @staticmethod
    def forward(ctx, features_rows, transform, queries_rows, keys_rows, frequencies, distance_transform, plan, csr, debug):
        indices, ranges, dists, densities, inv_total = csr
        lo, hi, P = plan.lo, plan.hi, plan.P
        if plan.world == 1 and lo == 0 and hi == P:
            feat, keys = features_rows.contiguous(), keys_rows.contiguous()
        else:
            # full-height staging; only the strip [jlo, jhi) is ever read by the kernels
            feat = torch.empty((P, features_rows.size(-1)), dtype=torch.float32, device=features_rows.device)
            keys = torch.empty((P, keys_rows.size(-1)), dtype=torch.float32, device=features_rows.device)
            feat[lo:hi] = features_rows
            keys[lo:hi] = keys_rows
            plan.gather([feat, keys])
        w, e, f, out = _C.aggregate_neighbors_rows(feat, transform, queries_rows, keys, frequencies, distance_transform,
                                                   indices, ranges, dists, densities, inv_total, (lo, hi), debug)
        ctx.plan, ctx.csr, ctx.debug = plan, csr, debug
        ctx.save_for_backward(feat, transform, queries_rows, keys, frequencies, distance_transform, w, e, f)
        return out

    @staticmethod
    def backward(ctx, g_rows):
        plan = ctx.plan
        indices, ranges, dists, densities, inv_total = ctx.csr
        feat, transform, queries_rows, keys, frequencies, distance_transform, w, e, f = ctx.saved_tensors
        lo, hi, P = plan.lo, plan.hi, plan.P
        g_feat = torch.empty_like(feat)
        g_keys = torch.empty_like(keys)
        if plan.dense:  # all-reduced as a whole
            g_feat.zero_()
            g_keys.zero_()
        else:           # the kernel adds into rows of the strip only, and only strip rows travel
            g_feat[plan.jlo:plan.jhi].zero_()
            g_keys[plan.jlo:plan.jhi].zero_()
        gt, gq, gfr, gdt = _C.aggregate_neighbors_rows_backward(
            feat, transform, queries_rows, keys, frequencies, distance_transform, indices, ranges, dists, densities, w, e, f,
            inv_total, g_rows.contiguous(), g_feat, g_keys, (lo, hi), ctx.debug)
        plan.scatter_add([g_feat, g_keys])
        if plan.world > 1:
            flat, recipe = pack_grads([gt, gfr, gdt])
            allreduce_grads(flat, plan.group)
            gt, gfr, gdt = unpack_grads(flat, recipe, 3)
        return g_feat[lo:hi], gt, gq, g_keys[lo:hi], gfr, gdt, None, None, None


class _AggregateRowsHeads(torch.autograd.Function):
    """All attention heads of one step in ONE node (model_pn.py:256-262 loops over ATTENTION_HEADS with the same
    features and frequencies): one halo exchange forward (features + every head's keys), one backward (the summed
    dL/dfeatures + every head's dL/dkeys) and one all-reduce of all shared-parameter gradients, instead of two exchanges
    and one all-reduce PER HEAD — at 8 ranks the per-collective latency, not the bytes, is what the aggregation waits on."""

    @staticmethod
    def forward(ctx, features_rows, frequencies, plan, csr, debug, H, *heads):
        indices, ranges, dists, densities, inv_total = csr
        lo, hi, P = plan.lo, plan.hi, plan.P
        T = heads[0::4]
        Q = [q.contiguous() for q in heads[1::4]]
        Kr = heads[2::4]
        DT = heads[3::4]
        local = plan.world == 1 and lo == 0 and hi == P
        if local:
            feat, keys = features_rows.contiguous(), [k.contiguous() for k in Kr]
        else:
            dev = features_rows.device
            feat = torch.empty((P, features_rows.size(-1)), dtype=torch.float32, device=dev)
            feat[lo:hi] = features_rows
            keys = []
            for k in Kr:
                kf = torch.empty((P, k.size(-1)), dtype=torch.float32, device=dev)
                kf[lo:hi] = k
                keys.append(kf)
            plan.gather([feat] + keys)
        outs, saved = [], []
        for h in range(H):
            w, e, f, out = _C.aggregate_neighbors_rows(feat, T[h], Q[h], keys[h], frequencies, DT[h], indices, ranges, dists,
                                                       densities, inv_total, (lo, hi), debug)
            outs.append(out)
            saved += [w, e, f]
        ctx.plan, ctx.csr, ctx.debug, ctx.H = plan, csr, debug, H
        ctx.save_for_backward(feat, frequencies, *T, *Q, *keys, *DT, *saved)
        ctx.set_materialize_grads(False)
        return tuple(outs)

    @staticmethod
    def backward(ctx, *g_outs):
        plan, H = ctx.plan, ctx.H
        indices, ranges, dists, densities, inv_total = ctx.csr
        sv = ctx.saved_tensors
        feat, frequencies = sv[0], sv[1]
        T, Q, keys, DT = sv[2:2 + H], sv[2 + H:2 + 2 * H], sv[2 + 2 * H:2 + 3 * H], sv[2 + 3 * H:2 + 4 * H]
        wef = sv[2 + 4 * H:]
        lo, hi, P = plan.lo, plan.hi, plan.P
        zero = (lambda t: t.zero_()) if plan.dense else (lambda t: t[plan.jlo:plan.jhi].zero_())
        g_feat = torch.empty_like(feat)
        zero(g_feat)  # every head ADDS its contribution into the same buffer
        g_keys, small, gq_all = [], [], []
        gfr_sum = None
        for h in range(H):
            gk = torch.empty_like(keys[h])
            zero(gk)
            g_keys.append(gk)
            if g_outs[h] is None:
                small += [torch.zeros_like(T[h]), torch.zeros_like(DT[h])]
                gq_all.append(torch.zeros_like(Q[h]))
                continue
            w, e, f = wef[3 * h:3 * h + 3]
            gt, gq, gfr, gdt = _C.aggregate_neighbors_rows_backward(
                feat, T[h], Q[h], keys[h], frequencies, DT[h], indices, ranges, dists, densities, w, e, f, inv_total,
                g_outs[h].contiguous(), g_feat, gk, (lo, hi), ctx.debug)
            gfr_sum = gfr if gfr_sum is None else gfr_sum + gfr
            small += [gt, gdt]
            gq_all.append(gq)
        if gfr_sum is None:
            gfr_sum = torch.zeros_like(frequencies)
        plan.scatter_add([g_feat] + g_keys)
        if plan.world > 1:
            flat, recipe = pack_grads([gfr_sum] + small)
            allreduce_grads(flat, plan.group)
            un = unpack_grads(flat, recipe, 1 + len(small))
            gfr_sum, small = un[0], un[1:]
        per_head = []
        for h in range(H):
            per_head += [small[2 * h], gq_all[h], g_keys[h][lo:hi], small[2 * h + 1]]
        return (g_feat[lo:hi], gfr_sum, None, None, None, None) + tuple(per_head)


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
        # a rank whose shard is empty still joins the bounds collective (with the neutral element) and then bins nothing
        if samples.size(0) == 0:
            mm = torch.cat([torch.full((D,), float("inf")), torch.full((D,), float("-inf"))]).to(means.device)
        else:
            mm = _C.sample_bounds(samples, D)
        allreduce_bounds(mm, D, self.group)
        super().preprocess(means, values, covariances, conics, samples, bounds=mm)

    def _rank_world(self):
        return _rank_world(self.group)

    def preprocess_aggregate(self, row_range=None, csr=None, dense_fraction=0.5):
        """Neighbour lists of this rank's contiguous range of query Gaussians (every Gaussian stays a candidate
        neighbour); the other rows are empty.  1/world of the CSR per GPU.  row_range = (lo, hi) overrides the even
        split, e.g. with this rank's window of balanced_ranges().  csr = (indices, ranges, dists, densities,
        inv_total_densities) installs caller-built lists for the window instead of searching (the reference API takes the
        CSR as plain inputs too, aggregate_neighbors.h:17-29).  dense_fraction: halo size (as a fraction of P) above
        which the row-parallel form exchanges whole windows with collectives instead of halo rows point to point."""
        rank, world = self._rank_world()
        P = self.means.size(0)
        self.row_range = shard_range(P, rank, world) if row_range is None else tuple(int(v) for v in row_range)
        if csr is None:
            csr = _C.preprocess_aggregate(self.means, self.conics, self.radii, self.debug, rows=self.row_range)
        self.indices, self.ranges, self.dists, self.densities, self.inv_total_densities = csr
        self.row_plan = RowHaloPlan(P, self.row_range, self.indices, self.group, dense_fraction=dense_fraction)
        self.row_windows = self.row_plan.windows

    def aggregate_neighbors(self, features, transform, queries, keys, frequencies, distance_transform):
        """Full [P, L] output on every rank; gradients of all six inputs summed over ranks."""
        return aggregate_neighbors(features, transform, queries, keys, frequencies, distance_transform, self.indices,
                                   self.ranges, self.dists, self.densities, self.inv_total_densities, self.debug,
                                   reduce=AggregateReduce(self.group, getattr(self, "row_windows", None)))

    def aggregate_neighbors_rows(self, features_rows, transform, queries_rows, keys_rows, frequencies, distance_transform):
        """Row-parallel form: features / queries / keys hold THIS rank's window rows [hi-lo, .] and so does the result;
        their gradients come back for the window's rows, the shared parameters' gradients summed over ranks."""
        csr = (self.indices, self.ranges, self.dists, self.densities, self.inv_total_densities)
        return _AggregateRows.apply(features_rows, transform, queries_rows, keys_rows, frequencies, distance_transform,
                                    self.row_plan, csr, self.debug)

    def aggregate_neighbors_rows_heads(self, features_rows, frequencies, heads):
        """All heads of a step at once.  heads = [(transform, queries_rows, keys_rows, distance_transform), ...] (the
        per-head arguments of aggregate_neighbors_rows; features and frequencies are shared, model_pn.py:256-262).
        Returns one [hi-lo, L] output per head.  Same results as calling aggregate_neighbors_rows per head, with one
        exchange each way and one all-reduce for all of them."""
        csr = (self.indices, self.ranges, self.dists, self.densities, self.inv_total_densities)
        flat = [t for hd in heads for t in hd]
        return list(_AggregateRowsHeads.apply(features_rows, frequencies, self.row_plan, csr, self.debug, len(heads), *flat))
