"""GPU parity of the neighbour-list build and neighbour aggregation (pigs_b200 through the reference-shaped API /
C ABI) against the CPU oracle, the committed reference-CUDA goldens and — when oracle/_ref/_C.so is present —
the live reference extension.  Bit-exact: indices, ranges.  1e-5 norm-wise: dists, densities, inv_total,
aggregation outputs and gradients (shared-parameter gradients vs the REFERENCE: AGG_SHARED_TOL, its own
fp32-atomic noise floor; vs the float64-accumulating oracle: 1e-5)."""
import os

import numpy as np
import pytest

from oracle import pigs_oracle as po
from tests import refcuda
from tests.fixtures import AGG_RECIPES, agg_params, covariances_from, synthetic, _pack
from tests.helpers import AGG_KEYS, AGG_SHARED_TOL, TOL, agg_upstream, oracle_state, rel_err, to_dev

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _prep(rec):
    from pigs_b200 import GaussianSampler
    d = to_dev(rec)
    s = GaussianSampler(False)
    s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
    s.preprocess_aggregate()
    return s, d


def _check_lists(s, want):
    idx, rng, dists, dens, inv = want
    np.testing.assert_array_equal(s.ranges.cpu().numpy(), rng)
    np.testing.assert_array_equal(s.indices.cpu().numpy(), idx)
    assert tuple(s.dists.shape) == dists.shape
    assert rel_err(s.dists.cpu().numpy(), dists) <= TOL
    assert rel_err(s.densities.cpu().numpy(), dens) <= TOL
    assert rel_err(s.inv_total_densities.cpu().numpy(), inv) <= TOL


@pytest.fixture(params=["cells", "rows", "mixed"])
def nb_path(request):
    """Neighbour search routes: 'cells' = the cell-cooperative kernels (shared id-sorted candidate list, the default),
    'rows' = every row through the per-row kernels (sort / bitmap), 'mixed' = coarse grid with a 256-candidate
    capacity, so dense cells overflow to the per-row kernels and sparse ones do not."""
    from pigs_b200 import _lib
    cap, div = {"cells": (4096, 2.0), "rows": (0, 2.0), "mixed": (256, 1.0)}[request.param]
    _lib.lib().pigs_set_neighbor_search(cap, div)
    yield request.param
    _lib.lib().pigs_set_neighbor_search(4096, 2.0)


@pytest.mark.parametrize("name", list(AGG_RECIPES))
def test_neighbor_lists_bit_exact_vs_oracle_and_golden(name, nb_path):
    rec = AGG_RECIPES[name]()
    st = oracle_state(rec)
    s, d = _prep(rec)
    _check_lists(s, po.neighbors(rec["means"], rec["conics"], st["radii"]))
    g = np.load(os.path.join(GOLD, "refcuda_%s.npz" % name))
    _check_lists(s, (g["indices"], g["ranges"], g["dists"], g["densities"], g["inv_total"]))


@pytest.fixture(params=["specialised", "generic"])
def agg_path(request):
    """L = K = 16, E = 25 recipes run aggregate16.cu by default; 'generic' forces aggregate.cu's kernels for them too."""
    from pigs_b200 import _lib
    _lib.lib().pigs_set_aggregate_specialised(1 if request.param == "specialised" else 0)
    yield request.param
    _lib.lib().pigs_set_aggregate_specialised(1)


@pytest.mark.parametrize("name", list(AGG_RECIPES))
def test_aggregate_forward_backward(name, agg_path):
    import torch
    rec = AGG_RECIPES[name]()
    st = oracle_state(rec)
    s, d = _prep(rec)
    idx, rng, dists, dens, inv = po.neighbors(rec["means"], rec["conics"], st["radii"])
    a = [rec[k] for k in AGG_KEYS]
    w, e, f, out = po.aggregate_forward(*a, idx, rng, dists, dens, inv)
    up = agg_upstream(rec)
    want = po.aggregate_backward(*a, idx, rng, dists, dens, w, e, f, inv, up)
    t = [d[k].clone().requires_grad_(True) for k in AGG_KEYS]
    got = s.aggregate_neighbors(*t)
    assert rel_err(got.detach().cpu().numpy(), out) <= TOL
    got.backward(torch.from_numpy(up).cuda())
    g = np.load(os.path.join(GOLD, "refcuda_%s.npz" % name))
    assert rel_err(got.detach().cpu().numpy(), g["out"]) <= TOL
    for x, wv, key in zip(t, want, AGG_KEYS):
        assert x.grad is not None, key
        assert rel_err(x.grad.cpu().numpy(), wv) <= TOL, key
        tol = AGG_SHARED_TOL if key in ("transform", "frequencies", "distance_transform") else TOL
        assert rel_err(x.grad.cpu().numpy(), g["grad_" + key]) <= tol, key
    # saved per-slot products (the reference returns them from the forward, aggregate_neighbors.cu:414)
    from pigs_b200 import _C
    wq, eq, fq, _ = _C.aggregate_neighbors(*[d[k] for k in AGG_KEYS], s.indices, s.ranges, s.dists, s.densities,
                                           s.inv_total_densities, False)
    for x, wv in ((wq, w), (eq, e), (fq, f)):
        assert rel_err(x.cpu().numpy(), wv) <= TOL


def _dense_cluster(P=2600, seed=3):
    """every Gaussian neighbours (almost) every other: rows beyond the shared-memory sort capacity -> bitmap path"""
    rng = np.random.default_rng(seed)
    means = rng.uniform(-0.2, 0.2, (P, 2))
    cov, con = covariances_from(np.exp(rng.uniform(-1.2, -0.8, (P, 2))), rng.normal(0, 0.2, P))
    return _pack(means, rng.uniform(-1, 1, (P, 1)), cov, con, means.copy())


@pytest.mark.parametrize("make", [_dense_cluster, lambda: synthetic(20000, 64, seed=9),
                                  lambda: synthetic(6000, 64, seed=10, edge_fraction=0.3, scale=1.0),
                                  lambda: synthetic(3000, 64, seed=11, D=1),
                                  lambda: synthetic(3000, 64, seed=12, scale=2.5)])
def test_neighbor_lists_larger_sets(make, nb_path):
    rec = make()
    rec["samples"] = rec["means"].copy()
    st = oracle_state(rec)
    s, d = _prep(rec)
    want = po.neighbors(rec["means"], rec["conics"], st["radii"])
    _check_lists(s, want)


def test_degenerate_rows_have_no_neighbours(nb_path):
    rec = AGG_RECIPES["agg_model"]()
    rec["covariances"][5] = 0.0   # det == 0 -> radius 0 -> skipped row and column (aggregate_neighbors.cu:29,38)
    rec["covariances"][77] = [0.25, 0.25, 0.25]
    st = oracle_state(rec)
    s, d = _prep(rec)
    want = po.neighbors(rec["means"], rec["conics"], st["radii"])
    _check_lists(s, want)
    assert float(s.inv_total_densities[5]) == pytest.approx(1e6, rel=1e-6)


@pytest.mark.parametrize("name", ["agg_model", "agg_model_1d", "agg_grid5"])
def test_vs_live_reference(name):
    import torch
    if not refcuda.available():
        pytest.skip("oracle/_ref/_C.so not built")
    rec = AGG_RECIPES[name]()
    s, d = _prep(rec)
    ref = refcuda.RefSampler(False)
    ref.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
    indices, ranges, dists, densities, inv_total = ref.preprocess_aggregate()
    assert torch.equal(s.indices, indices) and torch.equal(s.ranges, ranges)
    a = [d[k] for k in AGG_KEYS]
    w, e, f, out = ref.C.aggregate_neighbors(*a, indices, ranges, dists, densities, inv_total, False)
    got = s.aggregate_neighbors(*a)
    assert rel_err(got.cpu().numpy(), out.cpu().numpy()) <= TOL


def test_generic_sizes():
    """L, K, E away from the model's 16/16/25 (runtime-sized kernel): L=40, K=7, E=13."""
    import torch
    rec = synthetic(500, 64, seed=31)
    rec["samples"] = rec["means"].copy()
    rec.update(agg_params(500, 40, 7, 13, 2, seed=32, dt_rows=1))
    st = oracle_state(rec)
    s, d = _prep(rec)
    idx, rng, dists, dens, inv = po.neighbors(rec["means"], rec["conics"], st["radii"])
    a = [rec[k] for k in AGG_KEYS]
    w, e, f, out = po.aggregate_forward(*a, idx, rng, dists, dens, inv)
    up = agg_upstream(rec)
    want = po.aggregate_backward(*a, idx, rng, dists, dens, w, e, f, inv, up)
    t = [d[k].clone().requires_grad_(True) for k in AGG_KEYS]
    got = s.aggregate_neighbors(*t)
    assert rel_err(got.detach().cpu().numpy(), out) <= TOL
    got.backward(torch.from_numpy(up).cuda())
    for x, wv, key in zip(t, want, AGG_KEYS):
        assert rel_err(x.grad.cpu().numpy(), wv) <= TOL, key


@pytest.mark.parametrize("name", ["agg_model", "agg_model_1d", "agg_model_edge"])
def test_row_sharded_lists_and_aggregation(name, nb_path):
    """Multi-GPU aggregation without a second GPU: the CSR of each half of the query Gaussians
    holds exactly the rows of the full CSR, the two row-sharded forward outputs add up to the unsharded output bit for
    bit (disjoint rows), and the two backward contributions add up to the unsharded gradients."""
    import torch
    from pigs_b200 import _C
    rec = AGG_RECIPES[name]()
    s, d = _prep(rec)
    P = rec["means"].shape[0]
    full = (s.indices, s.ranges, s.dists, s.densities, s.inv_total_densities)
    a = [d[k] for k in AGG_KEYS]
    wf, ef, ff, out_full = _C.aggregate_neighbors(*a, *full, False)
    up = torch.from_numpy(agg_upstream(rec)).cuda()
    g_full = _C.aggregate_neighbors_backward(*a, s.indices, s.ranges, s.dists, s.densities, wf, ef, ff,
                                             s.inv_total_densities, up, False)
    starts_full = torch.cat([torch.zeros(1, dtype=torch.int64, device="cuda"), s.ranges[:-1]])
    out_sum = torch.zeros_like(out_full)
    g_sum = [torch.zeros_like(g) for g in g_full]
    cut = P // 2 + 1
    for lo, hi in ((0, cut), (cut, P)):
        idx, rng, dists, dens, inv = _C.preprocess_aggregate(s.means, s.conics, s.radii, False, rows=(lo, hi))
        cnt = torch.diff(rng, prepend=torch.zeros(1, dtype=torch.int64, device="cuda"))
        cnt_full = torch.diff(s.ranges, prepend=torch.zeros(1, dtype=torch.int64, device="cuda"))
        assert int(cnt[:lo].sum()) == 0 and int(cnt[hi:].sum()) == 0
        assert torch.equal(cnt[lo:hi], cnt_full[lo:hi])
        a0, b0 = int(starts_full[lo]), int(s.ranges[hi - 1])
        assert idx.numel() == b0 - a0
        assert torch.equal(idx, s.indices[a0:b0])          # the rank's rows are the full CSR's rows, bit for bit
        assert torch.equal(dists, s.dists[a0:b0])
        assert torch.equal(dens, s.densities[a0:b0])
        assert torch.equal(inv[lo:hi], s.inv_total_densities[lo:hi])
        w, e, f, out = _C.aggregate_neighbors(*a, idx, rng, dists, dens, inv, False)
        assert float(out[:lo].abs().sum()) == 0.0 and float(out[hi:].abs().sum()) == 0.0
        assert torch.equal(out[lo:hi], out_full[lo:hi])
        out_sum += out
        g = _C.aggregate_neighbors_backward(*a, idx, rng, dists, dens, w, e, f, inv, up, False)
        for acc, part in zip(g_sum, g):
            acc += part
    assert torch.equal(out_sum, out_full)
    for key, got, want in zip(AGG_KEYS, g_sum, g_full):
        assert rel_err(got.cpu().numpy(), want.cpu().numpy()) <= TOL, key


@pytest.mark.parametrize("name,specialised", [("agg_model", 1), ("agg_model", 0), ("agg_model_1d", 1), ("agg_grid5", 1)])
def test_row_windowed_kernels_touch_only_their_rows(name, specialised):
    """pigs_aggregate_forward_rows / _backward_rows (row-parallel multi-GPU aggregation): the window's rows equal the
    unwindowed result bit for bit, nothing outside the window is read or written (queries / upstream / output are
    window-sized tensors passed with a displaced base pointer, surrounded by guard rows), and the per-window gradient
    contributions add up to the unwindowed gradients."""
    import torch
    from pigs_b200 import _C, _lib
    rec = AGG_RECIPES[name]()
    s, d = _prep(rec)
    P = rec["means"].shape[0]
    a = [d[k] for k in AGG_KEYS]
    _lib.lib().pigs_set_aggregate_specialised(specialised)
    try:
        wf, ef, ff, out_full = _C.aggregate_neighbors(*a, s.indices, s.ranges, s.dists, s.densities, s.inv_total_densities, False)
        up = torch.from_numpy(agg_upstream(rec)).cuda()
        g_full = _C.aggregate_neighbors_backward(*a, s.indices, s.ranges, s.dists, s.densities, wf, ef, ff,
                                                 s.inv_total_densities, up, False)
        g_feat = torch.zeros_like(a[0])
        g_keys = torch.zeros_like(a[3])
        small = [torch.zeros_like(a[1]), torch.zeros_like(a[4]), torch.zeros_like(a[5])]
        gq_rows = []
        cuts = [0, P // 3, P // 3, (2 * P) // 3 + 1, P]  # includes an empty window
        for lo, hi in zip(cuts[:-1], cuts[1:]):
            idx, rng, dists, dens, inv = _C.preprocess_aggregate(s.means, s.conics, s.radii, False, rows=(lo, hi))
            K, L = a[2].shape[1], a[0].shape[1]
            guard = torch.full((hi - lo + 2, K), float("nan"), device="cuda")  # a NaN row either side of the window
            guard[1:-1] = a[2][lo:hi]
            q_rows = guard[1:-1]
            w, e, f, out = _C.aggregate_neighbors_rows(a[0], a[1], q_rows, a[3], a[4], a[5], idx, rng, dists, dens, inv,
                                                       (lo, hi), False)
            assert out.shape == (hi - lo, L)
            assert torch.equal(out, out_full[lo:hi])
            gt, gq, gfr, gdt = _C.aggregate_neighbors_rows_backward(a[0], a[1], q_rows, a[3], a[4], a[5], idx, rng, dists,
                                                                    dens, w, e, f, inv, up[lo:hi].contiguous(), g_feat,
                                                                    g_keys, (lo, hi), False)
            assert gq.shape == (hi - lo, K)
            gq_rows.append(gq)
            for acc, part in zip(small, (gt, gfr, gdt)):
                acc += part
        got = [g_feat, small[0], torch.cat(gq_rows, 0), g_keys, small[1], small[2]]
        for key, g, want in zip(AGG_KEYS, got, g_full):
            assert torch.isfinite(g).all(), key
            assert rel_err(g.cpu().numpy(), want.cpu().numpy()) <= TOL, key
    finally:
        _lib.lib().pigs_set_aggregate_specialised(1)


def test_multi_head_node_equals_per_head_calls():
    """ShardedGaussianSampler.aggregate_neighbors_rows_heads (all attention heads in one autograd node: one exchange each
    way) against one aggregate_neighbors call per head: same outputs, same gradients (features / frequencies summed over
    the heads).  Single process: the plan has no peers."""
    import torch
    from pigs_b200.sharded import ShardedGaussianSampler
    rec = AGG_RECIPES["agg_model"]()
    d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
    s = ShardedGaussianSampler(False)
    s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
    s.preprocess_aggregate()
    g = torch.Generator(device="cuda").manual_seed(4)
    H = 2
    def leaves():
        torch.manual_seed(7)
        feat = d["features"].clone().requires_grad_(True)
        freq = d["frequencies"].clone().requires_grad_(True)
        heads = [[(d[k] * (1.0 + 0.25 * h)).clone().requires_grad_(True) for k in ("transform", "queries", "keys", "distance_transform")]
                 for h in range(H)]
        return feat, freq, heads
    ups = [torch.randn(d["features"].shape, device="cuda", generator=g) for _ in range(H)]
    fa, qa, ha = leaves()
    outs_a = s.aggregate_neighbors_rows_heads(fa, qa, [tuple(h) for h in ha])
    torch.autograd.backward(outs_a, ups)
    fb, qb, hb = leaves()
    outs_b = [s.aggregate_neighbors(fb, h[0], h[1], h[2], qb, h[3]) for h in hb]
    torch.autograd.backward(outs_b, ups)
    for a, b in zip(outs_a, outs_b):
        assert torch.equal(a.detach(), b.detach())
    pairs = [(fa, fb, "features"), (qa, qb, "frequencies")] + [(x, y, "head%d.%d" % (h, i)) for h in range(H)
                                                               for i, (x, y) in enumerate(zip(ha[h], hb[h]))]
    for x, y, nm in pairs:
        assert x.grad is not None and y.grad is not None, nm
        assert rel_err(x.grad.cpu().numpy(), y.grad.cpu().numpy()) <= TOL, nm


def test_float64_gradcheck_like_the_reference_test():
    """The reference's one live test: torch.autograd.gradcheck of aggregate_neighbors on float64 tensors
    (test_neighbor_aggregation.py:15-27, 69-101: 5x5 grid, LATENT 2, KEY 4, EMBEDDING 21), which there needs the
    `typedef double FLOAT` build (cuda_sampler/config.h:20).  Here the generic kernels are instantiated for double."""
    import torch
    from pigs_b200 import GaussianSampler
    rec = AGG_RECIPES["agg_grid5"]()
    s, d = _prep(rec)
    g = torch.Generator().manual_seed(0)
    P = rec["means"].shape[0]
    mk = lambda *sh: torch.rand(sh, generator=g, dtype=torch.float64).cuda().requires_grad_(True)
    features, transform, queries, keys = mk(P, 2), mk(2, 2), mk(P, 4), mk(P, 4)
    frequencies = (torch.randn(5, generator=g, dtype=torch.float64) * 10).cuda().requires_grad_(True)
    distance_transform = mk(2, 42)
    out = s.aggregate_neighbors(features, transform, queries, keys, frequencies, distance_transform)
    assert out.dtype == torch.float64 and out.shape == (P, 2)
    # the float32 product path on the same values agrees with the double one to fp32 accuracy
    out32 = s.aggregate_neighbors(*[t.detach().float() for t in (features, transform, queries, keys, frequencies, distance_transform)])
    assert rel_err(out32.cpu().numpy(), out.detach().cpu().numpy()) <= TOL
    fn = lambda *a: s.aggregate_neighbors(*a)
    assert torch.autograd.gradcheck(fn, (features, transform, queries, keys, frequencies, distance_transform),
                                    nondet_tol=1e-9)  # double atomics: the summation order is not fixed (nor is the reference's)


def test_unaligned_feature_rows_take_the_generic_kernels():
    """A features tensor whose storage starts 4 bytes off a 16-byte boundary (a view into a larger buffer): the specialised
    kernels' 16-byte row loads / vector REDs do not apply, the call falls back to the generic kernels, same results."""
    import torch
    rec = AGG_RECIPES["agg_model"]()
    s, d = _prep(rec)
    t = [d[k].clone() for k in AGG_KEYS]
    want = s.aggregate_neighbors(*t)
    feat = t[0]
    buf = torch.empty(feat.numel() + 1, dtype=torch.float32, device="cuda")
    off = buf[1:].view(feat.shape)
    off.copy_(feat)
    assert off.data_ptr() % 16 == 4 and off.is_contiguous()
    off.requires_grad_(True)
    ref = feat.clone().requires_grad_(True)
    got = s.aggregate_neighbors(off, *t[1:])
    assert rel_err(got.detach().cpu().numpy(), want.cpu().numpy()) <= TOL
    up = torch.from_numpy(agg_upstream(rec)).cuda()
    got.backward(up)
    s.aggregate_neighbors(ref, *t[1:]).backward(up)
    assert rel_err(off.grad.cpu().numpy(), ref.grad.cpu().numpy()) <= TOL
