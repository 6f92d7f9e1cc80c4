"""Parity at BASELINE.json's FULL sizes (C2 = 100K x 1M, C4 = 500K anisotropic torus x 2M, C3 = 1M x 4M), where the
CPU oracle cannot evaluate the whole pair set (C3: 5.1e11 pairs):

  * binning bit-exact against the oracle on the whole input (the oracle's binning is O(P + N));
  * values / gradient / Hessian of a random SUBSET of the samples against the oracle run on that subset with the
    tile grid pinned to the full problem's bounds — per-sample results depend on the other samples only through
    the grid, so these rows must equal the full run's rows;
  * parameter gradients for an upstream gradient that is non-zero on that subset only, against the oracle's backward
    over the subset (the CUDA backward still walks the full pair set);
  * the adjoint identity <u(v), G> = <v, dL/dv(G)> over ALL samples (u is linear in the values), which ties the
    full-size forward to the full-size backward without any oracle.
"""
import numpy as np
import pytest

from oracle import pigs_oracle as po
from tests.fixtures import synthetic
from tests.helpers import TOL, ranges_np, rel_err

pytestmark = pytest.mark.gpu

CONFIGS = {
    "C2": (100_000, 1_000_000, {}, 600),
    "C4": (500_000, 2_000_000, dict(edge_fraction=0.1, anisotropy=4.0), 300),
    "C3": (1_000_000, 4_000_000, {}, 240),
}


@pytest.mark.timeout(900)
@pytest.mark.parametrize("name", list(CONFIGS))
def test_full_size_against_oracle_subset(name):
    import torch

    from pigs_b200 import GaussianSampler, _C
    P, N, kw, nsub = CONFIGS[name]
    rec = synthetic(P, N, seed=0, **kw)
    st = po.preprocess(rec["means"], rec["covariances"], rec["samples"])
    d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
    for k in ("means", "values", "conics"):
        d[k].requires_grad_(True)
    s = GaussianSampler(False)
    s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
    # ---- binning, whole input, bit-exact ----
    assert s.num_rendered == st["num_rendered"]
    np.testing.assert_array_equal(s.radii.cpu().numpy(), st["radii"])
    np.testing.assert_array_equal(ranges_np(s.ranges, st["blocks"]), st["ranges"])
    np.testing.assert_array_equal(ranges_np(s.sample_ranges, st["blocks"]), st["sample_ranges"])
    pl, _, sl = _C.point_lists(s.num_rendered, s.binning_buffer, s.sample_binning_buffer, N)
    np.testing.assert_array_equal(pl.cpu().numpy().view(np.uint32), st["point_list"])
    np.testing.assert_array_equal(sl.cpu().numpy().view(np.uint32), st["sample_list"])
    assert s.pair_count() == po.pair_count(st)

    # ---- forward on everything; a subset of rows against the oracle ----
    outs = s.sample_fused(7)
    rng = np.random.default_rng(5)
    sub = np.sort(rng.choice(N, nsub, replace=False))
    sub_samples = np.ascontiguousarray(rec["samples"][sub])
    st_sub = po.preprocess(rec["means"], rec["covariances"], sub_samples, bounds=(st["min_b"], st["max_b"]))
    np.testing.assert_array_equal(st_sub["grid"], st["grid"])
    ups_sub, full_ups = [], []
    for fn, got in enumerate(outs):
        want = po.forward(fn, st_sub, rec["means"], rec["values"], rec["conics"], sub_samples)
        e = rel_err(got.detach()[torch.from_numpy(sub).cuda()].cpu().numpy(), want)
        assert e <= TOL, (name, "forward", fn, e)
        g = rng.normal(size=want.shape).astype(np.float32)
        ups_sub.append(g)
        full = torch.zeros_like(got)
        full[torch.from_numpy(sub).cuda()] = torch.from_numpy(g).cuda()
        full_ups.append(full)

    # ---- backward: upstream non-zero on the subset only == oracle backward over the subset ----
    sum((o * g).sum() for o, g in zip(outs, full_ups)).backward()
    want = [np.zeros_like(rec[k], dtype=np.float64) for k in ("means", "values", "conics")]
    for fn in range(3):
        for acc, part in zip(want, po.backward(fn, st_sub, rec["means"], rec["values"], rec["conics"], sub_samples,
                                               ups_sub[fn])):
            acc += part.reshape(acc.shape)
    for k, w in zip(("means", "values", "conics"), want):
        e = rel_err(d[k].grad.cpu().numpy().reshape(w.shape), w)
        assert e <= TOL, (name, "backward", k, e)

    # ---- adjoint identity over ALL samples, per function ----
    nr, bb, sbb, rg, srg = s._sample_state
    for fn, o in enumerate(outs):
        G = torch.randn(o.shape, device="cuda", generator=torch.Generator(device="cuda").manual_seed(fn))
        grads = [None] * 4
        grads[fn] = G
        _, dv, _ = _C.sample_backward(1 << fn, d["means"].detach(), d["values"].detach(), d["conics"].detach(),
                                      d["samples"], nr, grads, bb, sbb, rg, srg, False)
        lhs = (o.detach().double() * G.double()).sum().item()
        rhs = (d["values"].detach().double() * dv.double()).sum().item()
        scale = (o.detach().double().abs() * G.double().abs()).sum().item()
        assert abs(lhs - rhs) <= 1e-5 * scale, (name, "adjoint", fn, lhs, rhs, scale)


@pytest.mark.timeout(900)
def test_neighbor_search_full_size_rows_against_oracle():
    """C5 part ii: the neighbour search at reference semantics at P = 1M (1.58e9 slots, 31.6 GB of
    CSR).  The oracle's all-pairs scan is O(P^2), so a random sample of rows is scanned against all P candidates on the
    CPU and compared with the same rows of the GPU lists: counts and indices bit-exact, dists / densities /
    inv_total to 1e-5; plus whole-list invariants (row totals, ascending ids, self in every row)."""
    import torch

    from pigs_b200 import GaussianSampler
    P = 1_000_000
    rec = synthetic(P, 1024, seed=0)
    d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
    s = GaussianSampler(False)
    s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
    radii = s.radii.cpu().numpy()   # bit-exact against the oracle in the test above (C3's Gaussians)
    s.preprocess_aggregate()
    ranges = s.ranges.cpu().numpy()
    M = int(ranges[-1])
    assert s.indices.numel() == M and M > 1_000_000_000
    rng = np.random.default_rng(11)
    rows = np.sort(rng.choice(P, 384, replace=False))
    counts, idx, dists, dens, inv = po.neighbor_rows(rec["means"], rec["conics"], radii, rows)
    assert counts.max() <= idx.shape[1]
    starts = np.concatenate([[0], ranges[:-1]])
    for k, i in enumerate(rows):
        a, b = int(starts[i]), int(ranges[i])
        c = int(counts[k])
        assert b - a == c, (i, b - a, c)
        np.testing.assert_array_equal(s.indices[a:b].cpu().numpy(), idx[k, :c])
        assert rel_err(s.dists[a:b].cpu().numpy(), dists[k, :c]) <= TOL
        assert rel_err(s.densities[a:b].cpu().numpy(), dens[k, :c]) <= TOL
    assert rel_err(s.inv_total_densities.cpu().numpy()[rows], inv) <= TOL
    # whole-list invariants, in chunks (the CSR is 12.6 GB of int64 indices)
    row_of = torch.repeat_interleave(torch.arange(P, device="cuda", dtype=torch.int32),
                                     torch.from_numpy(np.diff(ranges, prepend=0)).cuda())
    n_self = 0
    step = 1 << 28
    for a in range(0, M, step):
        b = min(M, a + step + 1)
        ind = s.indices[a:b]
        ro = row_of[a:b]
        same_row = ro[1:] == ro[:-1]
        asc = (ind[1:] > ind[:-1]) | (ind[1:] < 0) | (ind[:-1] < 0)
        assert bool((asc | ~same_row).all())
        n_self += int((ind[: min(step, M - a)] == ro[: min(step, M - a)]).sum())
    assert n_self == P   # every (valid) Gaussian neighbours itself exactly once


@pytest.mark.timeout(900)
def test_aggregation_full_size_c5_against_oracle():
    """C5 part i at BASELINE size: aggregate_neighbors forward + backward over a 16-regular CSR at
    P = 1M Gaussians, L = K = 16, E = 25 (model_pn.py:44-49) — the kernels of the bench's c5 leg — against the oracle
    on ALL rows (its loops are O(slots); OpenMP): output and the six gradients to 1e-5 norm-wise, plus a size-independent
    property: the operator is affine in the features."""
    import torch
    from scipy.spatial import cKDTree

    from pigs_b200 import GaussianSampler, _C
    from tests.fixtures import agg_params
    from tests.helpers import AGG_KEYS
    P, k = 1_000_000, 16
    rec = synthetic(P, 16, seed=0)
    d = {kk: torch.from_numpy(v).cuda() for kk, v in rec.items()}
    s = GaussianSampler(False)
    s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["means"])
    _, idx = cKDTree(rec["means"]).query(rec["means"], k=k, workers=-1)
    indices = torch.from_numpy(np.sort(idx.astype(np.int64), axis=1).reshape(-1)).cuda()
    ranges = (torch.arange(1, P + 1, device="cuda") * k).contiguous()
    i_of = torch.arange(P, device="cuda").repeat_interleave(k)
    X = d["means"][indices] - d["means"][i_of]
    X = torch.where(X.abs() > 1, X - 2 * torch.sign(X), X)
    dists = (X / (0.333 * s.radii[i_of] + 1e-6).unsqueeze(1)).contiguous()
    cj = d["conics"][indices]
    dens = torch.exp(-0.5 * (cj[:, 0] * X[:, 0] ** 2 + cj[:, 2] * X[:, 1] ** 2) - cj[:, 1] * X[:, 0] * X[:, 1]).contiguous()
    inv_total = (1.0 / (dens.view(P, k).sum(1) + 1e-6)).contiguous()
    del X, cj, i_of
    ap = agg_params(P, 16, 16, 25, 2, seed=100)
    a = [torch.from_numpy(ap[kk]).cuda().contiguous() for kk in AGG_KEYS]
    csr = (indices, ranges, dists, dens, inv_total)
    w, e, f, out = _C.aggregate_neighbors(*a, *csr, False)
    g = torch.randn(P, 16, device="cuda", generator=torch.Generator(device="cuda").manual_seed(9))
    grads = _C.aggregate_neighbors_backward(*a, indices, ranges, dists, dens, w, e, f, inv_total, g, False)
    # ---- the oracle on every row ----
    host = [ap[kk] for kk in AGG_KEYS]
    c_idx, c_rng, c_d, c_dn, c_inv = (t.cpu().numpy() for t in csr)
    ow, oe, of_, oout = po.aggregate_forward(*host, c_idx, c_rng, c_d, c_dn, c_inv)
    assert rel_err(out.cpu().numpy(), oout) <= TOL
    assert rel_err(w.cpu().numpy(), ow) <= TOL and rel_err(e.cpu().numpy(), oe) <= TOL and rel_err(f.cpu().numpy(), of_) <= TOL
    ograds = po.aggregate_backward(*host, c_idx, c_rng, c_d, c_dn, ow, oe, of_, c_inv, g.cpu().numpy())
    for key, got, want in zip(AGG_KEYS, grads, ograds):
        assert rel_err(got.cpu().numpy().reshape(want.shape), want) <= TOL, key
    # ---- the operator is AFFINE in the features (embedded = dw e + dw f feature, aggregate_neighbors.cu:190-200): an
    # affine combination (coefficients summing to 1) of feature sets maps to the same combination of outputs (no oracle) ----
    f2 = torch.randn_like(a[0])
    out2 = _C.aggregate_neighbors(f2, *a[1:], *csr, False)[3]
    mix = _C.aggregate_neighbors(1.75 * a[0] - 0.75 * f2, *a[1:], *csr, False)[3]
    lin = 1.75 * out - 0.75 * out2
    assert float((mix - lin).abs().max()) <= 1e-5 * float(lin.abs().max())
