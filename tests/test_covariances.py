"""Covariance / conic construction (gaussians.py:163-193).
CPU: the oracle restatement against goldens produced by the real gaussians.py + torch.autograd.
GPU (-m gpu): the fused CUDA kernels (through pigs_b200.gaussians -> C ABI) against the oracle and the goldens."""
import os

import numpy as np
import pytest

from oracle import pigs_oracle as po
from tests.helpers import TOL, rel_err

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "covariance_golden.npz")
CASES = ("model_pn", "derivatives", "wide")


def _case(name):
    z = np.load(GOLD)
    return {k.split("/", 1)[1]: z[k] for k in z.files if k.startswith(name + "/")}


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_golden(name):
    c = _case(name)
    cov, con = po.build_covariances(c["s"], c["t"])
    assert rel_err(cov, c["cov"]) <= TOL
    assert rel_err(con, c["con"]) <= TOL
    ds, dt = po.build_covariances_backward(c["s"], c["t"], c["g_cov"], c["g_con"])
    assert rel_err(ds, c["ds"]) <= TOL
    assert rel_err(dt, c["dt"]) <= TOL


def test_oracle_1d():
    s = np.array([[0.5], [2.0]], np.float32)
    cov, con = po.build_covariances(s, None)
    np.testing.assert_allclose(con, 1.0 / s)
    ds, dt = po.build_covariances_backward(s, None, np.ones_like(s), np.ones_like(s))
    np.testing.assert_allclose(ds, 1.0 - 1.0 / (s * s))


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_cuda_matches_oracle_and_golden(name):
    import torch

    from pigs_b200 import gaussians as pg
    c = _case(name)
    s = torch.from_numpy(c["s"]).cuda().requires_grad_(True)
    t = torch.from_numpy(c["t"]).cuda().requires_grad_(True)
    cov, con = pg.build_covariances(s, t)
    want_cov, want_con = po.build_covariances(c["s"], c["t"])
    assert rel_err(cov.detach().cpu().numpy(), want_cov) <= TOL
    assert rel_err(con.detach().cpu().numpy(), want_con) <= TOL
    assert rel_err(con.detach().cpu().numpy(), c["con"]) <= TOL
    g_cov, g_con = torch.from_numpy(c["g_cov"]).cuda(), torch.from_numpy(c["g_con"]).cuda()
    ((cov * g_cov).sum() + (con * g_con).sum()).backward()
    want_ds, want_dt = po.build_covariances_backward(c["s"], c["t"], c["g_cov"], c["g_con"])
    for got, want, gold in ((s.grad, want_ds, c["ds"]), (t.grad, want_dt, c["dt"])):
        assert rel_err(got.cpu().numpy(), want) <= TOL
        assert rel_err(got.cpu().numpy(), gold) <= TOL


@pytest.mark.gpu
def test_cuda_full_matrices_feed_the_sampler():
    """build_full_covariances / flatten_covariances round trip, gradients through only one output, and 1-D."""
    import torch

    from pigs_b200 import gaussians as pg
    c = _case("derivatives")
    s = torch.from_numpy(c["s"]).cuda().requires_grad_(True)
    t = torch.from_numpy(c["t"]).cuda().requires_grad_(True)
    full_cov, full_con = pg.build_full_covariances(s, t)
    assert full_cov.shape == (s.shape[0], 2, 2)
    assert torch.equal(full_cov[:, 0, 1], full_cov[:, 1, 0])
    cov, con = pg.flatten_covariances(full_cov, full_con)
    want_cov, want_con = po.build_covariances(c["s"], c["t"])
    assert rel_err(con.detach().cpu().numpy(), want_con) <= TOL
    g_con = torch.from_numpy(c["g_con"]).cuda()
    (con * g_con).sum().backward()  # covariance branch unused -> NULL upstream on the C ABI side
    want_ds, want_dt = po.build_covariances_backward(c["s"], c["t"], np.zeros_like(c["g_cov"]), c["g_con"])
    assert rel_err(s.grad.cpu().numpy(), want_ds) <= TOL
    assert rel_err(t.grad.cpu().numpy(), want_dt) <= TOL
    s1 = torch.tensor([[0.5], [2.0]], device="cuda", requires_grad=True)
    cov1, con1 = pg.build_covariances(s1, torch.zeros(2, 1, device="cuda"))
    assert torch.allclose(con1, 1.0 / s1)
    (cov1.sum() + con1.sum()).backward()
    assert torch.allclose(s1.grad, 1.0 - 1.0 / (s1 * s1))
