"""Split / prune as a device-side compaction (model_pn.py:578-610, 703-716;
test_no_mlp_1d.py:203-247).  CPU: the oracle restatement against goldens produced by the real model_pn.Model.split.
GPU (-m gpu): pigs_b200.densify (-> C ABI) against the oracle and the goldens.  The row ORDER of the two children of a
split Gaussian follows the sign of LAPACK's eigenvector in the reference, which the reference does not define; every
other row, and the two children as a set, are compared exactly / to 1e-5."""
import os

import numpy as np
import pytest

from oracle import pigs_oracle as po
from tests.helpers import TOL, rel_err

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "densify_golden.npz")
CASES = ("small", "c2", "all", "none")


def _case(name):
    z = np.load(GOLD)
    return {k.split("/", 1)[1]: z[k] for k in z.files if k.startswith(name + "/")}


def _check_against(c, out_means, out_scaling, out_transforms, out_u, n_keep):
    assert out_means.shape == c["out_means"].shape
    np.testing.assert_array_equal(out_means[:n_keep], c["out_means"][:n_keep])  # kept rows: plain copies
    np.testing.assert_array_equal(out_scaling, c["out_scaling"])
    np.testing.assert_array_equal(out_transforms, c["out_transforms"])
    np.testing.assert_array_equal(out_u, c["out_u"])
    if out_means.shape[0] > n_keep:
        assert rel_err(po.children_sorted(out_means, n_keep), po.children_sorted(c["out_means"], n_keep)) <= TOL


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_golden(name):
    c = _case(name)
    m, s, t, u, src, n_keep = po.split(c["means"], c["scaling"], c["transforms"], c["u"], c["cov"], c["indices"])
    assert n_keep == int((c["indices"] == 0).sum())
    _check_against(c, m, s, t, u, n_keep)


def test_oracle_plan_orders():
    keep = np.array([1, 0, 1, 1, 0], bool)
    ext = np.array([0, 0, 1, 0, 1], bool)
    src, nk, ne = po.compact_plan(5, keep, ext, 2)
    assert (nk, ne) == (3, 2) and src.tolist() == [0, 2, 3, 2, 2, 4, 4]
    src, nk, ne = po.compact_plan(5, keep, ext, 1)
    assert src.tolist() == [0, 2, 3, 2, 4]


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_cuda_split_matches_oracle_and_golden(name):
    import torch

    from pigs_b200 import densify
    c = _case(name)
    t = {k: torch.from_numpy(c[k]).cuda() for k in ("means", "scaling", "transforms", "u", "cov", "indices")}
    m, s, tr, u, src = densify.split(t["means"], t["scaling"], t["transforms"], t["u"], t["cov"], t["indices"])
    wm, ws, wt, wu, wsrc, n_keep = po.split(c["means"], c["scaling"], c["transforms"], c["u"], c["cov"], c["indices"])
    np.testing.assert_array_equal(src.cpu().numpy(), wsrc)
    got = [x.cpu().numpy() for x in (m, s, tr, u)]
    _check_against(c, *got, n_keep)
    np.testing.assert_array_equal(got[0][:n_keep], wm[:n_keep])
    if got[0].shape[0] > n_keep:
        assert rel_err(po.children_sorted(got[0], n_keep), po.children_sorted(wm, n_keep)) <= TOL


@pytest.mark.gpu
@pytest.mark.parametrize("m", [0, 1, 2, 3])
def test_cuda_plan_and_gather_match_oracle(m):
    import torch

    from pigs_b200 import densify
    rng = np.random.default_rng(5 + m)
    P = 100_003
    keep, ext = rng.random(P) < 0.9, rng.random(P) < 0.1
    x = rng.normal(size=(P, 3)).astype(np.float32)
    src, nk, ne = densify.plan(P, torch.from_numpy(keep).cuda(), torch.from_numpy(ext).cuda(), m)
    wsrc, wnk, wne = po.compact_plan(P, keep, ext, m)
    assert (nk, ne) == (wnk, wne)
    np.testing.assert_array_equal(src.cpu().numpy(), wsrc)
    xd = torch.from_numpy(x).cuda()
    np.testing.assert_array_equal(densify.gather_rows(xd, src, nk).cpu().numpy(), x[wsrc])
    z = densify.gather_rows(xd, src, nk, appended_scale=0.0).cpu().numpy()  # Adam moments of the clones
    np.testing.assert_array_equal(z[:nk], x[wsrc[:nk]])
    assert not z[nk:].any()


@pytest.mark.gpu
def test_cuda_prune_only_and_1d():
    import torch

    from pigs_b200 import densify
    rng = np.random.default_rng(9)
    P = 5000
    keep = rng.random(P) < 0.5
    src, nk, ne = densify.plan(P, keep=torch.from_numpy(keep).cuda())
    assert ne == 0 and src.cpu().numpy().tolist() == np.nonzero(keep)[0].tolist()
    means = rng.uniform(-1, 1, (P, 1)).astype(np.float32)
    cov = np.exp(rng.normal(size=(P, 1)) - 4).astype(np.float32)
    u = rng.normal(size=(P, 1)).astype(np.float32)
    idx = rng.random(P) < 0.1
    td = lambda a: torch.from_numpy(a).cuda()
    m, s, t, uu, src = densify.split(td(means), td(cov), td(cov), td(u), td(cov), td(idx))
    wm, _, _, wu, wsrc, n_keep = po.split(means, cov, cov, u, cov, idx)
    np.testing.assert_array_equal(src.cpu().numpy(), wsrc)
    assert rel_err(m.cpu().numpy(), wm) <= TOL  # 1-D: v = +1 on both sides, the order is defined
    np.testing.assert_array_equal(uu.cpu().numpy(), wu)


@pytest.mark.gpu
def test_split_is_differentiable_like_the_reference_chain():
    """Model.split (model_pn.py:578-605) indexes, halves and concatenates NON-LEAF tensors (the dynamics network's outputs),
    so autograd must flow through it: gradients of pigs_b200.densify.split equal those of the torch formulation
    (x[~idx], x[idx].repeat_interleave(2) (/ 2), cat; displacement constant) for every input."""
    import torch
    from pigs_b200 import densify
    g = torch.Generator().manual_seed(3)
    P = 257
    base = {"means": torch.rand(P, 2, generator=g) * 2 - 1, "scaling": torch.rand(P, 2, generator=g) * 0.01 + 0.001,
            "transforms": torch.randn(P, 1, generator=g) * 0.3, "u": torch.rand(P, 2, generator=g) * 2 - 1}
    idx = (torch.arange(P) % 5 == 0).cuda()
    s0, s1, t = base["scaling"][:, 0], base["scaling"][:, 1], torch.tanh(base["transforms"][:, 0]) * torch.sqrt(base["scaling"][:, 0] * base["scaling"][:, 1])
    cov = torch.stack([s0, t, s1], -1).cuda()
    # two graphs over the same non-leaf inputs (x = leaf * 1.5, as the network's `means + dmeans * mask` would be)
    leaves_a = {k: v.clone().cuda().requires_grad_(True) for k, v in base.items()}
    leaves_b = {k: v.clone().cuda().requires_grad_(True) for k, v in base.items()}
    xa = {k: v * 1.5 for k, v in leaves_a.items()}
    xb = {k: v * 1.5 for k, v in leaves_b.items()}
    om, os_, ot, ou, src = densify.split(xa["means"], xa["scaling"], xa["transforms"], xa["u"], cov, idx)
    assert all(o.grad_fn is not None for o in (om, os_, ot, ou))
    n_keep = int((~idx).sum())
    # torch formulation; the (constant) displacement is taken from our forward so both sides place the children alike
    disp = (om[n_keep:] - xa["means"].detach()[src[n_keep:]]).detach()
    rm = torch.cat((xb["means"][~idx], xb["means"][idx].repeat_interleave(2, 0) + disp), 0)
    rs = torch.cat((xb["scaling"][~idx], xb["scaling"][idx].repeat_interleave(2, 0)), 0)
    rt = torch.cat((xb["transforms"][~idx], xb["transforms"][idx].repeat_interleave(2, 0)), 0)
    ru = torch.cat((xb["u"][~idx], xb["u"][idx].repeat_interleave(2, 0) / 2.0), 0)
    assert float((om.detach() - rm.detach()).abs().max()) <= 1e-6  # m + (child - m) re-rounds the child
    for a, b in ((os_, rs), (ot, rt), (ou, ru)):
        assert torch.equal(a.detach(), b.detach())
    w = [torch.randn(o.shape, generator=torch.Generator().manual_seed(9 + i)).cuda() for i, o in enumerate((om, os_, ot, ou))]
    sum((o * wi).sum() for o, wi in zip((om, os_, ot, ou), w)).backward()
    sum((o * wi).sum() for o, wi in zip((rm, rs, rt, ru), w)).backward()
    for k in base:
        assert leaves_a[k].grad is not None
        assert rel_err(leaves_a[k].grad.cpu().numpy(), leaves_b[k].grad.cpu().numpy()) <= 1e-6, k
    # masks and integer rows: torch indexing, appended rows set like model_pn.py:602-605
    mask = (torch.arange(P) % 3 == 0).cuda().reshape(P, 1)
    mo = densify.gather_rows(mask, src, n_keep, appended_value=True)
    assert mo.dtype == torch.bool and mo.shape == (src.numel(), 1) and bool(mo[n_keep:].all())
    assert torch.equal(mo[:n_keep], mask[~idx])
    bo = densify.gather_rows(torch.ones(P, 1).cuda(), src, n_keep, appended_value=0.0)
    assert float(bo[n_keep:].abs().sum()) == 0.0 and float(bo[:n_keep].sum()) == n_keep
