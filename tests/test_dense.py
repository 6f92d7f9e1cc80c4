"""Dense all-pairs path (gaussians.py:48-58 sample_gaussians, :89-101 gaussian_derivative,
:103-116 gaussian_derivative2): FULL [n, d, d] conics, x = sample - mean, no cutoff, no torus.
CPU: the torch port used as the timed CPU baseline (oracle/dense_torch.py) against goldens produced by importing the
real gaussians.py (tests/golden/make_dense_golden.py -> dense_full_golden.npz), incl. 1-D and non-symmetric conics.
GPU (-m gpu): pigs_b200.gaussians' CUDA implementation (pigs_dense_forward / _backward through the C ABI) against the
same goldens and, at a size where both operands are chunked over the grid, against the port."""
import os

import numpy as np
import pytest

from tests.helpers import TOL, rel_err

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "dense_full_golden.npz")
CASES = ("f2", "f2c3", "f2asym", "f1", "f1c2")
FNS = (("u", "sample_gaussians"), ("du", "gaussian_derivative"), ("ddu", "gaussian_derivative2"))


def _case(tag):
    z = np.load(GOLD)
    return {k[len(tag) + 1:]: z[k] for k in z.files if k.startswith(tag + "_")}


@pytest.mark.parametrize("tag", CASES)
def test_port_matches_reference_golden(tag):
    import torch

    from oracle import dense_torch as dt
    c = _case(tag)
    for key, name in FNS:
        m, k, v = (torch.from_numpy(c[x]).clone().requires_grad_(True) for x in ("means", "conics", "values"))
        o = getattr(dt, name)(m, k, v, torch.from_numpy(c["samples"]))
        assert rel_err(o.detach().numpy(), c[key]) <= TOL, (tag, key)
        gm, gk, gv = torch.autograd.grad((o * torch.from_numpy(c["g" + key])).sum(), (m, k, v))
        for got, suffix in ((gm, "_dmeans"), (gk, "_dconics"), (gv, "_dvalues")):
            assert rel_err(got.numpy(), c[key + suffix]) <= TOL, (tag, key, suffix)


@pytest.mark.gpu
@pytest.mark.parametrize("tag", CASES)
def test_dense_cuda_matches_reference_golden(tag):
    import torch

    from pigs_b200 import gaussians as pg
    c = _case(tag)
    samples = torch.from_numpy(c["samples"]).cuda()
    for key, name in FNS:
        m, k, v = (torch.from_numpy(c[x]).cuda().requires_grad_(True) for x in ("means", "conics", "values"))
        o = getattr(pg, name)(m, k, v, samples)
        assert tuple(o.shape) == c[key].shape
        assert rel_err(o.detach().cpu().numpy(), c[key]) <= TOL, (tag, key)
        (o * torch.from_numpy(c["g" + key]).cuda()).sum().backward()
        for t, suffix in ((m, "_dmeans"), (k, "_dconics"), (v, "_dvalues")):
            assert rel_err(t.grad.cpu().numpy(), c[key + suffix]) <= TOL, (tag, key, suffix)
    # the three functions in ONE launch, gradients of all three upstreams in one backward
    m, k, v = (torch.from_numpy(c[x]).cuda().requires_grad_(True) for x in ("means", "conics", "values"))
    outs = pg.dense_sample(7, m, k, v, samples)
    sum((o * torch.from_numpy(c["g" + key]).cuda()).sum() for o, (key, _) in zip(outs, FNS)).backward()
    for o, (key, _) in zip(outs, FNS):
        assert rel_err(o.detach().cpu().numpy(), c[key]) <= TOL
    for t, suffix in ((m, "_dmeans"), (k, "_dconics"), (v, "_dvalues")):
        want = sum(c[key + suffix].astype(np.float64) for key, _ in FNS)
        assert rel_err(t.grad.cpu().numpy(), want) <= TOL, (tag, "fused", suffix)


@pytest.mark.gpu
@pytest.mark.parametrize("D,C,n,N", [(2, 1, 5000, 3000), (1, 2, 6000, 2500)])
def test_dense_cuda_chunked_grid_matches_port(D, C, n, N):
    """Sizes where the forward cuts the Gaussians and the backward cuts the samples into blockIdx.y chunks (RED path)."""
    import torch

    from oracle import dense_torch as dt
    from pigs_b200 import gaussians as pg
    g = torch.Generator().manual_seed(5)
    means = torch.rand(n, D, generator=g) * 2 - 1
    s = torch.exp(torch.randn(n, D, generator=g) * 0.3 - 4.0)
    if D == 2:
        off = torch.tanh(torch.randn(n, generator=g) * 0.3) * torch.sqrt(s[:, 0] * s[:, 1])
        cov = torch.stack([s[:, 0], off, off, s[:, 1]], -1).reshape(n, 2, 2)
        con = torch.linalg.inv(cov)
    else:
        con = (1.0 / s).reshape(n, 1, 1)
    values = torch.rand(n, C, generator=g) * 2 - 1
    samples = torch.rand(N, D, generator=g) * 2 - 1
    ups = [torch.randn(N, C, generator=g), torch.randn(N, D, C, generator=g), torch.randn(N, D, D, C, generator=g)]
    m, k, v = (x.clone().requires_grad_(True) for x in (means, con, values))
    want = [dt.sample_gaussians(m, k, v, samples), dt.gaussian_derivative(m, k, v, samples),
            dt.gaussian_derivative2(m, k, v, samples)]
    sum((o * u).sum() for o, u in zip(want, ups)).backward()
    mc, kc, vc = (x.cuda().requires_grad_(True) for x in (means, con, values))
    outs = pg.dense_sample(7, mc, kc, vc, samples.cuda())
    sum((o * u.cuda()).sum() for o, u in zip(outs, ups)).backward()
    for o, w in zip(outs, want):
        assert rel_err(o.detach().cpu().numpy(), w.detach().numpy()) <= TOL
    for a, b in ((mc, m), (kc, k), (vc, v)):
        assert rel_err(a.grad.cpu().numpy(), b.grad.numpy()) <= 2 * TOL


@pytest.mark.gpu
@pytest.mark.parametrize("tag", ["f2", "f2asym", "f1", "f1c2"])
def test_dense_cuda_multi_slot_kernels_match_golden(tag):
    """The 4-/2-samples-per-thread forward and 2-Gaussians-per-thread backward instantiations (picked by size in
    production) forced at golden size."""
    import torch

    from pigs_b200 import _lib
    from pigs_b200 import gaussians as pg
    c = _case(tag)
    L = _lib.lib()
    L.pigs_set_dense_slots(4, 2)
    try:
        m, k, v = (torch.from_numpy(c[x]).cuda().requires_grad_(True) for x in ("means", "conics", "values"))
        outs = pg.dense_sample(7, m, k, v, torch.from_numpy(c["samples"]).cuda())
        sum((o * torch.from_numpy(c["g" + key]).cuda()).sum() for o, (key, _) in zip(outs, FNS)).backward()
        for o, (key, _) in zip(outs, FNS):
            assert rel_err(o.detach().cpu().numpy(), c[key]) <= TOL, key
        for t, suffix in ((m, "_dmeans"), (k, "_dconics"), (v, "_dvalues")):
            want = sum(c[key + suffix].astype(np.float64) for key, _ in FNS)
            assert rel_err(t.grad.cpu().numpy(), want) <= TOL, suffix
    finally:
        L.pigs_set_dense_slots(0, 0)


@pytest.mark.gpu
def test_dense_cuda_rejects_bad_shapes():
    import torch

    from pigs_b200 import gaussians as pg
    m, v, x = torch.zeros(4, 2).cuda(), torch.zeros(4, 1).cuda(), torch.zeros(8, 2).cuda()
    with pytest.raises(RuntimeError):
        pg.sample_gaussians(m, torch.zeros(4, 3).cuda(), v, x)  # flattened conics are not the dense path's input
    with pytest.raises(RuntimeError):
        pg.sample_gaussians(torch.zeros(4, 3).cuda(), torch.zeros(4, 3, 3).cuda(), v, torch.zeros(8, 3).cuda())


@pytest.mark.gpu
def test_dense_cuda_empty_operands():
    import torch

    from pigs_b200 import gaussians as pg
    m, k, v = torch.zeros(0, 2).cuda(), torch.zeros(0, 2, 2).cuda(), torch.zeros(0, 1).cuda()
    u, ux, uxx = pg.dense_sample(7, m, k, v, torch.rand(5, 2).cuda())
    assert u.shape == (5, 1) and ux.shape == (5, 2, 1) and uxx.shape == (5, 2, 2, 1) and not u.any() and not uxx.any()
    m, k, v = torch.rand(3, 2).cuda().requires_grad_(True), torch.eye(2).repeat(3, 1, 1).cuda(), torch.ones(3, 1).cuda()
    (u,) = pg.dense_sample(1, m, k, v, torch.zeros(0, 2).cuda())
    assert u.shape == (0, 1)
    u.sum().backward()
    assert m.grad is not None and not m.grad.any()


@pytest.mark.gpu
def test_dense_cuda_large_rows_and_adjoint():
    """20K Gaussians x 400K points (8e9 pairs): the sizes where the 8-sample packed forward and the 2-Gaussian backward run
    on grids cut into several whole waves of chunks (csrc/dense.cu dn_chunks).  Rows: u / grad u / Hessian of a sample
    subset against the PyTorch formulation on the host (every row depends on its own sample only).  Whole problem, no
    oracle: the operator is linear in the values, so <u(v), G> = <v, dL/dv(G)> ties the forward to the backward, and the
    gradient w.r.t. the values of a sum over a sample subset equals the port's."""
    import torch

    from oracle import dense_torch as dt
    from pigs_b200 import gaussians as pg
    n, N, D = 20_000, 400_000, 2
    g = torch.Generator().manual_seed(17)
    means = torch.rand(n, D, generator=g) * 2 - 1
    s = torch.exp(torch.randn(n, D, generator=g) * 0.3 - 5.0)
    off = torch.tanh(torch.randn(n, generator=g) * 0.3) * torch.sqrt(s[:, 0] * s[:, 1])
    con = torch.linalg.inv(torch.stack([s[:, 0], off, off, s[:, 1]], -1).reshape(n, 2, 2))
    values = torch.rand(n, 1, generator=g) * 2 - 1
    samples = torch.rand(N, D, generator=g) * 2 - 1
    mc, kc, vc, sc = means.cuda(), con.cuda(), values.cuda().requires_grad_(True), samples.cuda()
    u, ux, uxx = pg.dense_sample(7, mc, kc, vc, sc)
    rows = torch.arange(0, N, 1571)
    sub = samples[rows]
    for got, want in ((u, dt.sample_gaussians(means, con, values, sub)), (ux, dt.gaussian_derivative(means, con, values, sub)),
                      (uxx, dt.gaussian_derivative2(means, con, values, sub))):
        assert rel_err(got.detach()[rows.cuda()].cpu().numpy(), want.numpy()) <= TOL
    G = torch.randn(N, 1, device="cuda", generator=torch.Generator(device="cuda").manual_seed(3))
    (dv,) = torch.autograd.grad((u * G).sum(), vc, retain_graph=True)
    lhs = (u.detach().double() * G.double()).sum().item()
    rhs = (vc.detach().double() * dv.double()).sum().item()
    assert abs(lhs - rhs) <= 1e-5 * max(abs(lhs), abs(rhs), 1.0)
    # gradients of a loss that touches the subset's rows only: equal to the port's over the subset
    Gs = torch.zeros(N, D, 1, device="cuda")
    Gs[rows.cuda()] = torch.randn(rows.numel(), D, 1, device="cuda", generator=torch.Generator(device="cuda").manual_seed(4))
    mg, kg = mc.clone().requires_grad_(True), kc.clone().requires_grad_(True)
    ux2 = pg.gaussian_derivative(mg, kg, vc.detach(), sc)
    (ux2 * Gs).sum().backward()
    mh, kh = means.clone().requires_grad_(True), con.clone().requires_grad_(True)
    (dt.gaussian_derivative(mh, kh, values, sub) * Gs[rows.cuda()].cpu()).sum().backward()
    assert rel_err(mg.grad.cpu().numpy(), mh.grad.numpy()) <= 2 * TOL
    assert rel_err(kg.grad.cpu().numpy(), kh.grad.numpy()) <= 2 * TOL
