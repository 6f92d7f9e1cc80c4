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
