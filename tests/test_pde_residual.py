"""Laplacian output + PDE right-hand side / residual loss (model_pn.py:612-625, 822-837).
CPU: the oracle restatement against goldens produced by the real model_pn.Model.pde_rhs + torch.autograd
(tests/golden/make_pde_golden.py).
GPU (-m gpu): PIGS_FN_LAPL through GaussianSampler.sample_laplacian() against the trace of the oracle's Hessian
(forward and backward, alone and inside the fused plan), and the fused residual kernels (pigs_b200.pde -> C ABI)
against the oracle and the goldens."""
import os

import numpy as np
import pytest

from oracle import pigs_oracle as po
from tests.fixtures import synthetic
from tests.helpers import TOL, oracle_state, rel_err, upstream

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "pde_golden.npz")
CASES = ("diffusion", "diffusion_c3", "burgers", "burgers_c2", "wave")


def _case(name):
    z = np.load(GOLD)
    return {k.split("/", 1)[1]: z[k] for k in z.files if k.startswith(name + "/")}


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_golden(name):
    c = _case(name)
    prob, nu, dt = int(c["problem"]), float(c["nu"]), float(c["dt"])
    r, loss = po.pde_residual_loss(prob, c["u"], c["ux"], c["lap"], c["ut"], nu, dt)
    assert rel_err(r, c["residual"]) <= TOL
    assert abs(loss - float(c["loss"])) <= TOL * abs(float(c["loss"]))
    d_u, d_ux, d_lap, d_ut = po.pde_residual_backward(prob, c["u"], c["ux"], c["lap"], c["ut"], nu, dt, float(c["up"]))
    assert rel_err(d_lap, c["d_lap"]) <= TOL
    assert rel_err(d_ut, c["d_ut"]) <= TOL
    assert rel_err(d_u, c["d_u"]) <= TOL
    assert rel_err(d_ux, c["d_ux"]) <= TOL


NS_CASES = ("navier_stokes", "navier_stokes_b")


@pytest.mark.parametrize("name", NS_CASES)
def test_oracle_navier_stokes_matches_reference_golden(name):
    """Vorticity extraction + NAVIER_STOKES residual / loss / gradients against the real model_pn chain
    (model_pn.py:779-781 -> pde_rhs :629-631 -> :843-845)."""
    c = _case(name)
    nu, dt, up = float(c["nu"]), float(c["dt"]), float(c["up"])
    np.testing.assert_array_equal(po.vorticity(c["ux"]), c["w"])
    np.testing.assert_array_equal(po.vorticity(c["uxx"]), c["wx"])
    np.testing.assert_array_equal(po.vorticity(c["uxxx"]), c["wxx"])
    np.testing.assert_array_equal(po.vorticity(c["ux"]) - po.vorticity(c["ux_prev"]), c["wt"])
    r, div, loss = po.ns_residual_loss(c["u"], c["ux"], c["wx"], c["wxx"], c["wt"], nu, dt)
    assert rel_err(r, c["residual"]) <= TOL
    assert rel_err(div, c["divergence"]) <= TOL
    assert abs(loss - float(c["loss"])) <= TOL * abs(float(c["loss"]))
    d_u, d_ux, d_wx, d_wxx, d_wt = po.ns_residual_backward(c["u"], c["ux"], c["wx"], c["wxx"], c["wt"], nu, dt, up)
    for got, key in ((d_u, "d_u"), (d_ux, "d_ux"), (d_wx, "d_wx"), (d_wxx, "d_wxx"), (d_wt, "d_wt")):
        assert rel_err(got, c[key]) <= TOL, key
    # the reference's gradients w.r.t. the sampler outputs are the vorticity adjoint of those w.r.t. wx / wxx
    assert rel_err(po.vorticity_backward(d_wx), c["d_uxx"]) <= TOL
    assert rel_err(po.vorticity_backward(d_wxx), c["d_uxxx"]) <= TOL


LEVEL_CASES = ("levels_diffusion", "levels_burgers", "levels_burgers_c2", "levels_burgers_forward",
               "levels_burgers_backward", "levels_wave")
LEVEL_KEYS = ("u1", "u0", "ux1", "ux0", "lap1", "lap0")


def _ts_of(c):
    rule = int(c["rule"])  # 2 = TRAPEZOID (per-sample ts), 0 = FORWARD (x[-2]), 1 = BACKWARD (x[-1])
    return c["ts"] if rule == 2 else float(rule)


@pytest.mark.parametrize("name", LEVEL_CASES)
def test_oracle_timelevels_match_reference_golden(name):
    """compute_loss's blend -> ut -> dt * pde_rhs -> loss chain (model_pn.py:794-837) with the real pde_rhs + autograd."""
    c = _case(name)
    prob, nu, dt, up, ts = int(c["problem"]), float(c["nu"]), float(c["dt"]), float(c["up"]), _ts_of(c)
    u, ux, lap = (c["u1"], c["u0"]), (c["ux1"], c["ux0"]), (c["lap1"], c["lap0"])
    r, loss = po.pde_residual_loss_timelevels(prob, ts, u, ux, lap, nu, dt)
    assert rel_err(r, c["residual"]) <= TOL
    assert abs(loss - float(c["loss"])) <= TOL * abs(float(c["loss"]))
    grads = po.pde_residual_backward_timelevels(prob, ts, u, ux, lap, nu, dt, up)
    for got, key in zip(grads, LEVEL_KEYS):
        assert rel_err(got, c["d_" + key]) <= TOL, key


NS_LEVEL_KEYS = ("u1", "u0", "ux1", "ux0", "wx1", "wx0", "wxx1", "wxx0", "w1", "w0")


def test_oracle_navier_stokes_timelevels_match_reference_golden():
    c = _case("levels_navier_stokes")
    nu, dt, up = float(c["nu"]), float(c["dt"]), float(c["up"])
    pairs = [(c[NS_LEVEL_KEYS[i]], c[NS_LEVEL_KEYS[i + 1]]) for i in range(0, 10, 2)]
    r, div, loss = po.ns_residual_loss_timelevels(c["ts"], *pairs, nu, dt)
    np.testing.assert_array_equal(r, c["residual"])      # same fp32 operation order as the torch lines
    np.testing.assert_array_equal(div, c["divergence"])
    assert abs(loss - float(c["loss"])) <= TOL * abs(float(c["loss"]))
    for got, key in zip(po.ns_residual_backward_timelevels(c["ts"], *pairs, nu, dt, up), NS_LEVEL_KEYS):
        assert rel_err(got, c["d_" + key]) <= TOL, key


def test_problem_enum_matches_reference_values():
    from pigs_b200._C import PDE_BURGERS, PDE_DIFFUSION, PDE_WAVE  # import of the glue only; no library call
    from pigs_b200.pde import Problem
    assert (Problem.DIFFUSION.value, Problem.BURGERS.value, Problem.WAVE.value) == (PDE_DIFFUSION, PDE_BURGERS, PDE_WAVE)
    assert (po.PDE_DIFFUSION, po.PDE_BURGERS, po.PDE_WAVE) == (PDE_DIFFUSION, PDE_BURGERS, PDE_WAVE)
    for name in CASES:  # the goldens carry the value of the real model_pn.Problem member
        want = {"diffusion": PDE_DIFFUSION, "burgers": PDE_BURGERS, "wave": PDE_WAVE}[name.split("_")[0]]
        assert int(_case(name)["problem"]) == want


def test_oracle_laplacian_is_trace_and_adjoint():
    rec = synthetic(300, 2000, seed=5)
    st = oracle_state(rec)
    a = (rec["means"], rec["values"], rec["conics"], rec["samples"])
    H = po.forward(po.FN_HESS, st, *a)
    lap = po.laplacian(st, *a)
    np.testing.assert_array_equal(lap, H[:, 0, 0] + H[:, 1, 1])
    g = np.random.default_rng(0).normal(size=lap.shape).astype(np.float32)
    dm, dv, dc = po.laplacian_backward(st, *a, g)
    # <dlap/dv . g> == <lap, g> because lap is linear in the values
    assert abs((dv * rec["values"]).sum() - (lap.astype(np.float64) * g).sum()) <= 1e-4 * np.abs(lap * g).sum()


# ------------------------------------------------------------------------------------------------ GPU
def _sampler(rec, **kw):
    import torch

    from diff_gaussian_sampling import GaussianSampler
    d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
    for k in ("means", "values", "conics"):
        d[k].requires_grad_(True)
    s = GaussianSampler(False, **kw)
    s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
    return s, d


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [dict(P=3000, N=20000), dict(P=800, N=5000, C=2), dict(P=500, N=3000, D=1),
                                   dict(P=2000, N=9000, edge_fraction=0.2, anisotropy=4.0)])
def test_laplacian_forward_backward_matches_oracle(shape):
    import torch
    kw = dict(shape)
    rec = synthetic(kw.pop("P"), kw.pop("N"), seed=11, **kw)
    st = oracle_state(rec)
    a = (rec["means"], rec["values"], rec["conics"], rec["samples"])
    s, d = _sampler(rec)
    lap = s.sample_laplacian()
    want = po.laplacian(st, *a)
    assert lap.shape == want.shape
    assert rel_err(lap.detach().cpu().numpy(), want) <= TOL
    g = upstream(rec, 0, seed=3).reshape(want.shape)
    (lap * torch.from_numpy(g).cuda()).sum().backward()
    for k, w in zip(("means", "values", "conics"), po.laplacian_backward(st, *a, g)):
        assert rel_err(d[k].grad.cpu().numpy().reshape(w.shape), w) <= TOL, k


@pytest.mark.gpu
def test_laplacian_in_fused_plan_equals_unfused():
    """u + grad u + Laplacian requested as three calls: the second preprocess cycle runs them as ONE kernel."""
    import torch
    rec = synthetic(4000, 30000, seed=12)
    st = oracle_state(rec)
    a = (rec["means"], rec["values"], rec["conics"], rec["samples"])
    s, d = _sampler(rec, fused=True)
    for cycle in range(2):
        if cycle:
            s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
        u, ux, lap = s.sample_gaussians(), s.sample_gaussians_derivative(), s.sample_laplacian()
    assert u.grad_fn is ux.grad_fn is lap.grad_fn  # one autograd node -> one fused backward
    assert rel_err(u.detach().cpu().numpy(), po.forward(0, st, *a)) <= TOL
    assert rel_err(ux.detach().cpu().numpy(), po.forward(1, st, *a)) <= TOL
    assert rel_err(lap.detach().cpu().numpy(), po.laplacian(st, *a)) <= TOL
    gs = [upstream(rec, 0), upstream(rec, 1), upstream(rec, 0, seed=9)]
    (u * torch.from_numpy(gs[0]).cuda()).sum().add((ux * torch.from_numpy(gs[1]).cuda()).sum()).add(
        (lap * torch.from_numpy(gs[2]).cuda()).sum()).backward()
    want = [np.zeros(rec[k].shape, np.float64) for k in ("means", "values", "conics")]
    parts = (po.backward(0, st, *a, gs[0]), po.backward(1, st, *a, gs[1]), po.laplacian_backward(st, *a, gs[2]))
    for part in parts:
        for acc, p in zip(want, part):
            acc += p.reshape(acc.shape)
    for k, w in zip(("means", "values", "conics"), want):
        assert rel_err(d[k].grad.cpu().numpy().reshape(w.shape), w) <= TOL, k


@pytest.mark.gpu
def test_laplacian_and_hessian_in_one_cycle_run_unfused():
    rec = synthetic(1000, 6000, seed=13)
    st = oracle_state(rec)
    a = (rec["means"], rec["values"], rec["conics"], rec["samples"])
    s, d = _sampler(rec, fused=True)
    for cycle in range(2):
        if cycle:
            s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
        H, lap = s.sample_gaussians_laplacian(), s.sample_laplacian()
    assert rel_err(H.detach().cpu().numpy(), po.forward(2, st, *a)) <= TOL
    assert rel_err(lap.detach().cpu().numpy(), po.laplacian(st, *a)) <= TOL


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_residual_cuda_matches_oracle_and_golden(name):
    import torch

    from pigs_b200 import pde
    c = _case(name)
    prob, nu, dt = int(c["problem"]), float(c["nu"]), float(c["dt"])
    t = {k: torch.from_numpy(c[k]).cuda().requires_grad_(True) for k in ("u", "ux", "lap", "ut")}
    loss = pde.pde_residual_loss(pde.Problem(prob), t["u"], t["ux"], t["lap"], t["ut"], dt, nu)
    r_want, loss_want = po.pde_residual_loss(prob, c["u"], c["ux"], c["lap"], c["ut"], nu, dt)
    assert abs(float(loss.detach()) - loss_want) <= TOL * abs(loss_want)
    assert abs(float(loss.detach()) - float(c["loss"])) <= TOL * abs(float(c["loss"]))
    res, loss2 = pde.pde_residual(pde.Problem(prob), t["u"].detach(), t["ux"].detach(), t["lap"].detach(),
                                  t["ut"].detach(), dt, nu)
    assert float(loss2) == float(loss.detach())  # deterministic two-stage reduction
    assert rel_err(res.cpu().numpy(), r_want) <= TOL
    assert rel_err(res.cpu().numpy(), c["residual"]) <= TOL
    (loss * float(c["up"])).backward()
    want = po.pde_residual_backward(prob, c["u"], c["ux"], c["lap"], c["ut"], nu, dt, float(c["up"]))
    for k, w in zip(("u", "ux", "lap", "ut"), want):
        got = t[k].grad
        got = np.zeros(c[k].shape, np.float32) if got is None else got.cpu().numpy()
        assert rel_err(got, w) <= TOL, k
        assert rel_err(got, c["d_" + k]) <= TOL, k


@pytest.mark.gpu
@pytest.mark.parametrize("name", NS_CASES)
def test_navier_stokes_cuda_matches_oracle_and_golden(name):
    """pigs_b200.pde.vorticity + navier_stokes_residual_loss (C ABI) chained like Model.sample / compute_loss, forward and
    backward to the sampler's outputs, against the oracle and the goldens of the real model_pn chain."""
    import torch

    from pigs_b200 import pde
    c = _case(name)
    nu, dt, up = float(c["nu"]), float(c["dt"]), float(c["up"])
    t = {k: torch.from_numpy(c[k]).cuda().requires_grad_(True) for k in ("u", "ux", "uxx", "uxxx", "wt")}
    w, wx, wxx = pde.vorticity(t["ux"]), pde.vorticity(t["uxx"]), pde.vorticity(t["uxxx"])
    np.testing.assert_array_equal(w.detach().cpu().numpy(), c["w"])        # one fp32 subtraction: bit-exact
    np.testing.assert_array_equal(wx.detach().cpu().numpy(), c["wx"])
    np.testing.assert_array_equal(wxx.detach().cpu().numpy(), c["wxx"])
    loss = pde.navier_stokes_residual_loss(t["u"], t["ux"], wx, wxx, t["wt"], dt, nu)
    r_want, div_want, loss_want = po.ns_residual_loss(c["u"], c["ux"], c["wx"], c["wxx"], c["wt"], nu, dt)
    assert abs(float(loss.detach()) - loss_want) <= TOL * abs(loss_want)
    assert abs(float(loss.detach()) - float(c["loss"])) <= TOL * abs(float(c["loss"]))
    res, div, loss2 = pde.navier_stokes_residual(t["u"].detach(), t["ux"].detach(), wx.detach(), wxx.detach(),
                                                 t["wt"].detach(), dt, nu)
    assert float(loss2) == float(loss.detach())  # deterministic two-stage reduction
    np.testing.assert_array_equal(res.cpu().numpy(), r_want)   # same fp32 operation order as the torch expressions
    np.testing.assert_array_equal(div.cpu().numpy(), div_want)
    assert rel_err(res.cpu().numpy(), c["residual"]) <= TOL
    (loss * up).backward()
    for k in ("u", "ux", "uxx", "uxxx", "wt"):
        assert rel_err(t[k].grad.cpu().numpy(), c["d_" + k]) <= TOL, k
    want = po.ns_residual_backward(c["u"], c["ux"], c["wx"], c["wxx"], c["wt"], nu, dt, up)
    assert rel_err(t["u"].grad.cpu().numpy(), want[0]) <= TOL
    assert rel_err(t["uxx"].grad.cpu().numpy(), po.vorticity_backward(want[2])) <= TOL
    assert rel_err(t["uxxx"].grad.cpu().numpy(), po.vorticity_backward(want[3])) <= TOL


@pytest.mark.gpu
@pytest.mark.parametrize("name", LEVEL_CASES)
def test_timelevels_cuda_matches_oracle_and_golden(name):
    """Time integration inside the residual kernels (pigs_pde_residual_timelevels_*) against the oracle and the goldens of
    the real compute_loss chain, all three integration rules."""
    import torch

    from pigs_b200 import pde
    c = _case(name)
    prob, nu, dt, up, ts = int(c["problem"]), float(c["nu"]), float(c["dt"]), float(c["up"]), _ts_of(c)
    t = {k: torch.from_numpy(c[k]).cuda().requires_grad_(True) for k in LEVEL_KEYS}
    ts_dev = torch.from_numpy(ts).cuda() if isinstance(ts, np.ndarray) else ts
    loss = pde.pde_residual_loss_timelevels(pde.Problem(prob), ts_dev, (t["u1"], t["u0"]), (t["ux1"], t["ux0"]),
                                            (t["lap1"], t["lap0"]), dt, nu)
    u, ux, lap = (c["u1"], c["u0"]), (c["ux1"], c["ux0"]), (c["lap1"], c["lap0"])
    _, loss_want = po.pde_residual_loss_timelevels(prob, ts, u, ux, lap, nu, dt)
    assert abs(float(loss.detach()) - loss_want) <= TOL * abs(loss_want)
    assert abs(float(loss.detach()) - float(c["loss"])) <= TOL * abs(float(c["loss"]))
    (loss * up).backward()
    want = po.pde_residual_backward_timelevels(prob, ts, u, ux, lap, nu, dt, up)
    for k, w in zip(LEVEL_KEYS, want):
        got = t[k].grad
        got = np.zeros(c[k].shape, np.float32) if got is None else got.cpu().numpy()
        assert rel_err(got, w) <= TOL, k
        assert rel_err(got, c["d_" + k]) <= TOL, k
    # the [N, C] residual itself
    from pigs_b200 import _C
    _, res = _C.pde_residual_timelevels_forward(prob, ts_dev, (t["u1"].detach(), t["u0"].detach()),
                                                (t["ux1"].detach(), t["ux0"].detach()),
                                                (t["lap1"].detach(), t["lap0"].detach()), nu, dt, want_residual=True)
    assert rel_err(res.cpu().numpy(), c["residual"]) <= TOL


@pytest.mark.gpu
def test_navier_stokes_timelevels_cuda_matches_oracle_and_golden():
    import torch

    from pigs_b200 import _C, pde
    c = _case("levels_navier_stokes")
    nu, dt, up = float(c["nu"]), float(c["dt"]), float(c["up"])
    t = {k: torch.from_numpy(c[k]).cuda().requires_grad_(True) for k in NS_LEVEL_KEYS}
    ts = torch.from_numpy(c["ts"]).cuda()
    pairs = [(t[NS_LEVEL_KEYS[i]], t[NS_LEVEL_KEYS[i + 1]]) for i in range(0, 10, 2)]
    loss = pde.navier_stokes_residual_loss_timelevels(ts, *pairs, dt, nu)
    assert abs(float(loss.detach()) - float(c["loss"])) <= TOL * abs(float(c["loss"]))
    _, res, div = _C.pde_ns_residual_timelevels_forward(ts, *[(a.detach(), b.detach()) for a, b in pairs], nu, dt,
                                                        want_residual=True)
    np.testing.assert_array_equal(res.cpu().numpy(), c["residual"])     # fp32 op order of the torch lines
    np.testing.assert_array_equal(div.cpu().numpy(), c["divergence"])
    (loss * up).backward()
    np_pairs = [(c[NS_LEVEL_KEYS[i]], c[NS_LEVEL_KEYS[i + 1]]) for i in range(0, 10, 2)]
    want = po.ns_residual_backward_timelevels(c["ts"], *np_pairs, nu, dt, up)
    for k, w in zip(NS_LEVEL_KEYS, want):
        assert rel_err(t[k].grad.cpu().numpy(), w) <= TOL, k
        assert rel_err(t[k].grad.cpu().numpy(), c["d_" + k]) <= TOL, k


@pytest.mark.gpu
def test_navier_stokes_chain_through_sampler():
    """Two-channel Gaussians -> u, grad u, Hessian, third derivative from the sampler -> vorticity terms -> fused
    NAVIER_STOKES residual loss -> backward to means / values / conics, against the oracle chain."""
    import torch

    from pigs_b200 import pde
    rec = synthetic(600, 4000, seed=21, C=2)
    st = oracle_state(rec)
    a = (rec["means"], rec["values"], rec["conics"], rec["samples"])
    N = rec["samples"].shape[0]
    nu, dt = 0.05, 0.1
    wt = (0.1 * np.random.default_rng(4).normal(size=N)).astype(np.float32)
    s, d = _sampler(rec)
    u, ux, uxx, uxxx = (s.sample_gaussians(), s.sample_gaussians_derivative(), s.sample_gaussians_laplacian(),
                        s.sample_gaussians_third_derivative())
    loss = pde.navier_stokes_residual_loss(u, ux, pde.vorticity(uxx), pde.vorticity(uxxx), torch.from_numpy(wt).cuda(),
                                           dt, nu)
    loss.backward()
    o = [po.forward(fn, st, *a) for fn in range(4)]
    wxo, wxxo = po.vorticity(o[2]), po.vorticity(o[3])
    _, _, loss_want = po.ns_residual_loss(o[0], o[1], wxo, wxxo, wt, nu, dt)
    assert abs(float(loss.detach()) - loss_want) <= TOL * abs(loss_want)
    g_u, g_ux, g_wx, g_wxx, _ = po.ns_residual_backward(o[0], o[1], wxo, wxxo, wt, nu, dt, 1.0)
    ups = (g_u, g_ux, po.vorticity_backward(g_wx), po.vorticity_backward(g_wxx))
    want = [np.zeros(rec[k].shape, np.float64) for k in ("means", "values", "conics")]
    for fn, g in enumerate(ups):
        for acc, p in zip(want, po.backward(fn, st, *a, g.astype(np.float32).reshape(o[fn].shape))):
            acc += p.reshape(acc.shape)
    for k, w in zip(("means", "values", "conics"), want):
        assert rel_err(d[k].grad.cpu().numpy().reshape(w.shape), w) <= 2 * TOL, k


@pytest.mark.gpu
def test_residual_large_and_chain_through_sampler():
    """End to end like compute_loss: sampler (u, grad u, Laplacian fused) -> fused Burgers residual loss -> backward
    to the Gaussian parameters, against the oracle chain."""
    import torch

    from pigs_b200 import pde
    rec = synthetic(3000, 50000, seed=14)
    st = oracle_state(rec)
    a = (rec["means"], rec["values"], rec["conics"], rec["samples"])
    nu, dt = 0.02, 0.05
    ut = (0.1 * np.random.default_rng(3).normal(size=(rec["samples"].shape[0], 1))).astype(np.float32)
    s, d = _sampler(rec)
    u, ux, lap = s.sample_fused(1 | 2 | 16)
    loss = pde.pde_residual_loss(pde.Problem.BURGERS, u, ux, lap, torch.from_numpy(ut).cuda(), dt, nu)
    loss.backward()
    uo, uxo, lo = po.forward(0, st, *a), po.forward(1, st, *a), po.laplacian(st, *a)
    _, loss_want = po.pde_residual_loss(po.PDE_BURGERS, uo, uxo, lo, ut, nu, dt)
    assert abs(float(loss.detach()) - loss_want) <= TOL * abs(loss_want)
    g_u, g_ux, g_lap, _ = po.pde_residual_backward(po.PDE_BURGERS, uo, uxo, lo, ut, nu, dt, 1.0)
    want = [np.zeros(rec[k].shape, np.float64) for k in ("means", "values", "conics")]
    parts = (po.backward(0, st, *a, g_u.astype(np.float32)), po.backward(1, st, *a, g_ux.astype(np.float32)),
             po.laplacian_backward(st, *a, g_lap.astype(np.float32)))
    for part in parts:
        for acc, p in zip(want, part):
            acc += p.reshape(acc.shape)
    for k, w in zip(("means", "values", "conics"), want):
        assert rel_err(d[k].grad.cpu().numpy().reshape(w.shape), w) <= 2 * TOL, k


@pytest.mark.gpu
def test_navier_stokes_inputs_at_odd_float_offsets():
    """Views that start 4 bytes off a 16-byte boundary are copied by the binding (the kernels read whole rows as vectors;
    the C ABI itself answers a misaligned pointer with 'bad argument')."""
    import torch

    from pigs_b200 import _lib, pde
    c = _case("navier_stokes_b")
    nu, dt = float(c["nu"]), float(c["dt"])

    def off(a):
        t = torch.from_numpy(a).cuda()
        buf = torch.empty(t.numel() + 1, dtype=torch.float32, device="cuda")
        v = buf[1:].view(t.shape)
        v.copy_(t)
        return v

    uxx, uxxx = off(c["uxx"]), off(c["uxxx"])
    assert uxx.data_ptr() % 16 == 4
    np.testing.assert_array_equal(pde.vorticity(uxx).cpu().numpy(), c["wx"])
    loss = pde.navier_stokes_residual_loss(off(c["u"]), off(c["ux"]), pde.vorticity(uxx), pde.vorticity(uxxx), off(c["wt"]), dt, nu)
    assert abs(float(loss) - float(c["loss"])) <= TOL * abs(float(c["loss"]))
    L = _lib.lib()
    out = torch.empty(c["wx"].size, dtype=torch.float32, device="cuda")
    import ctypes
    rc = L.pigs_vorticity_forward(out.numel(), ctypes.c_void_p(uxx.data_ptr()), ctypes.c_void_p(out.data_ptr()), None)
    assert rc != 0 and b"bad argument" in L.pigs_last_error()
