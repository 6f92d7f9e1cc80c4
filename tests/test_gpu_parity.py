"""GPU parity tests: the CUDA product (through pigs_b200's reference-shaped Python API, i.e. the C ABI of
libpigs_b200.so) against the CPU oracle on seeded inputs.  Bit-exact for binning, 1e-5 norm-wise for values,
derivatives and parameter gradients (BASELINE.json north_star)."""
import numpy as np
import pytest

from tests.fixtures import RECIPES, synthetic
from tests.helpers import TOL, oracle_state, ranges_np, rel_err, to_dev, upstream
from oracle import pigs_oracle as po

pytestmark = pytest.mark.gpu

NAMES = list(RECIPES)


@pytest.fixture(scope="module")
def torch_mod():
    import torch
    assert torch.cuda.is_available()
    return torch


def _sampler(rec, fused=False):
    from pigs_b200 import GaussianSampler
    d = to_dev(rec)
    s = GaussianSampler(False, fused=fused)
    s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
    return s, d


@pytest.mark.parametrize("name", NAMES)
def test_binning_bit_exact(name, torch_mod):
    from pigs_b200 import _C
    rec = RECIPES[name]()
    st = oracle_state(rec)
    s, d = _sampler(rec)
    N = rec["samples"].shape[0]
    assert s.num_rendered == st["num_rendered"]
    assert s.ranges.numel() == st["blocks"] * 8 + 8
    np.testing.assert_array_equal(s.radii.cpu().numpy().view(np.uint32), st["radii"].view(np.uint32))
    np.testing.assert_array_equal(ranges_np(s.ranges, st["blocks"]), st["ranges"])
    np.testing.assert_array_equal(ranges_np(s.sample_ranges, st["blocks"]), st["sample_ranges"])
    pl, pt, sl = _C.point_lists(s.num_rendered, s.binning_buffer, s.sample_binning_buffer, N)
    np.testing.assert_array_equal(pl.cpu().numpy().view(np.uint32), st["point_list"])
    np.testing.assert_array_equal(sl.cpu().numpy().view(np.uint32), st["sample_list"])
    assert s.pair_count() == po.pair_count(st)


@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("fn", [0, 1, 2, 3])
def test_forward_matches_oracle(name, fn, torch_mod):
    rec = RECIPES[name]()
    st = oracle_state(rec)
    want = po.forward(fn, st, rec["means"], rec["values"], rec["conics"], rec["samples"])
    s, d = _sampler(rec)
    got = [s.sample_gaussians, s.sample_gaussians_derivative, s.sample_gaussians_laplacian,
           s.sample_gaussians_third_derivative][fn]()
    assert tuple(got.shape) == want.shape
    assert rel_err(got.cpu().numpy(), want) <= TOL


@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("fn", [0, 1, 2, 3])
def test_backward_matches_oracle(name, fn, torch_mod):
    torch = torch_mod
    rec = RECIPES[name]()
    st = oracle_state(rec)
    g = upstream(rec, fn)
    want = po.backward(fn, st, rec["means"], rec["values"], rec["conics"], rec["samples"], g)
    d = to_dev(rec)
    for k in ("means", "values", "conics"):
        d[k].requires_grad_(True)
    from pigs_b200 import GaussianSampler
    s = GaussianSampler(False, fused=False)
    s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
    out = [s.sample_gaussians, s.sample_gaussians_derivative, s.sample_gaussians_laplacian,
           s.sample_gaussians_third_derivative][fn]()
    out.backward(torch.from_numpy(g).cuda())
    for got, w, nm in zip((d["means"].grad, d["values"].grad, d["conics"].grad), want, ("means", "values", "conics")):
        assert got is not None, nm
        assert rel_err(got.cpu().numpy().reshape(w.shape), w) <= TOL, nm


@pytest.mark.parametrize("name", ["synth2d", "synth2d_c2", "synth1d", "edge_aniso", "wide", "synth2d_c3"])
def test_fused_equals_separate(name, torch_mod):
    """value+gradient+Hessian in ONE kernel (mask 7) vs the three single-function kernels, fwd and bwd."""
    torch = torch_mod
    rec = RECIPES[name]()
    st = oracle_state(rec)
    d = to_dev(rec)
    for k in ("means", "values", "conics"):
        d[k].requires_grad_(True)
    from pigs_b200 import GaussianSampler
    s = GaussianSampler(False, fused=False)
    s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
    u, du, ddu = s.sample_fused(7)
    gs = [upstream(rec, fn) for fn in range(3)]
    for fn, got in enumerate((u, du, ddu)):
        want = po.forward(fn, st, rec["means"], rec["values"], rec["conics"], rec["samples"])
        assert rel_err(got.detach().cpu().numpy(), want) <= TOL
    loss = sum((o * torch.from_numpy(g).cuda()).sum() for o, g in zip((u, du, ddu), gs))
    loss.backward()
    want = [np.zeros_like(x, dtype=np.float64) for x in (rec["means"], rec["values"], rec["conics"])]
    for fn in range(3):
        for acc, part in zip(want, po.backward(fn, st, rec["means"], rec["values"], rec["conics"], rec["samples"], gs[fn])):
            acc += part.reshape(acc.shape)
    for got, w, nm in zip((d["means"].grad, d["values"].grad, d["conics"].grad), want, ("means", "values", "conics")):
        assert rel_err(got.cpu().numpy().reshape(w.shape), w) <= TOL, nm


def test_adaptive_fusion_and_separate_backward(torch_mod):
    """fused='auto': second preprocess cycle evaluates the remembered set in one launch; each output can still be
    back-propagated by its own backward() call (the reference has three independent autograd nodes)."""
    torch = torch_mod
    from pigs_b200 import GaussianSampler, _lib
    rec = RECIPES["synth2d"]()
    d = to_dev(rec)
    for k in ("means", "values", "conics"):
        d[k].requires_grad_(True)
    s = GaussianSampler(False, fused=True)
    ref = {}
    for cycle in range(2):
        s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
        n0 = _lib.lib().pigs_launch_count()
        u = s.sample_gaussians()
        du = s.sample_gaussians_derivative()
        ddu = s.sample_gaussians_laplacian()
        launches = _lib.lib().pigs_launch_count() - n0
        outs = dict(u=u, du=du, ddu=ddu)
        if cycle == 0:
            ref = {k: v.detach().clone() for k, v in outs.items()}
            first_launches = launches
        else:
            assert launches < first_launches  # one fused evaluation instead of three
            for k, v in outs.items():
                assert rel_err(v.detach().cpu().numpy(), ref[k].cpu().numpy()) <= 2e-6
            u.sum().backward()
            g1 = d["means"].grad.clone()
            du.sum().backward()
            assert not torch.equal(g1, d["means"].grad)


def test_empty_inputs(torch_mod):
    torch = torch_mod
    from pigs_b200 import GaussianSampler
    s = GaussianSampler(False)
    z = lambda *shape: torch.zeros(*shape, device="cuda")
    s.preprocess(z(0, 2), z(0, 1), z(0, 3), z(0, 3), z(10, 2))
    assert s.num_rendered == 0 and tuple(s.sample_gaussians().shape) == (10, 1)
    s.preprocess(z(4, 2), z(4, 1), z(4, 3), z(4, 3), z(0, 2))
    assert tuple(s.sample_gaussians_laplacian().shape) == (0, 2, 2, 1)


def test_degenerate_gaussians_are_dropped(torch_mod):
    """det == 0 covariance -> radius 0, no tile keys (forward.cu:55-57); NaN-free outputs."""
    torch = torch_mod
    rec = RECIPES["derivatives"]()
    rec["covariances"][3] = [0.25, 0.25, 0.25]  # exact products: det == 0
    rec["covariances"][7] = 0.0
    st = oracle_state(rec)
    s, d = _sampler(rec)
    assert s.num_rendered == st["num_rendered"]
    r = s.radii.cpu().numpy()
    assert r[3] == 0 and r[7] == 0
    want = po.forward(0, st, rec["means"], rec["values"], rec["conics"], rec["samples"])
    assert rel_err(s.sample_gaussians().cpu().numpy(), want) <= TOL


def test_noncontiguous_and_dtype_errors(torch_mod):
    torch = torch_mod
    from pigs_b200 import GaussianSampler
    rec = RECIPES["derivatives"]()
    d = to_dev(rec)
    s = GaussianSampler(False)
    big = torch.zeros(d["samples"].shape[0], 4, device="cuda")
    big[:, :2] = d["samples"]
    s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], big[:, :2])  # silently .contiguous()'d
    a = s.sample_gaussians()
    s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
    b = s.sample_gaussians()  # same inputs; the default mode's RED.ADD order varies run to run
    assert float((a - b).abs().max()) <= 2e-6 * float(b.abs().max())
    with pytest.raises(RuntimeError):
        s.preprocess(d["means"].double(), d["values"], d["covariances"], d["conics"], d["samples"])


@pytest.mark.gpu
def test_inputs_at_odd_float_offsets(torch_mod):
    """Every input a contiguous view that starts 4 bytes off an 8-/16-byte boundary (a slice of a larger buffer): no
    kernel may assume more alignment of the caller's tensors than a float's.  Same bins, same values, same gradients."""
    torch = torch_mod
    from pigs_b200 import GaussianSampler, _C
    rec = synthetic(1500, 9000, seed=77)

    def off(a):
        t = torch.from_numpy(a).cuda()
        buf = torch.empty(t.numel() + 1, dtype=t.dtype, device="cuda")
        v = buf[1:].view(t.shape)
        v.copy_(t)
        assert v.data_ptr() % 8 == 4 and v.is_contiguous()
        return v

    res = []
    for make in (lambda a: torch.from_numpy(a).cuda(), off):
        d = {k: make(v) for k, v in rec.items()}
        for k in ("means", "values", "conics"):
            d[k].requires_grad_(True)
        s = GaussianSampler(False)
        s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
        outs = s.sample_fused(7)
        sum((o * torch.from_numpy(upstream(rec, fn)).cuda()).sum() for fn, o in enumerate(outs)).backward()
        pl, _, sl = _C.point_lists(s.num_rendered, s.binning_buffer, s.sample_binning_buffer, rec["samples"].shape[0])
        res.append(dict(nr=s.num_rendered, pl=pl.cpu().numpy(), sl=sl.cpu().numpy(), radii=s.radii.cpu().numpy(),
                        outs=[o.detach().cpu().numpy() for o in outs],
                        grads=[d[k].grad.cpu().numpy() for k in ("means", "values", "conics")]))
    a, b = res
    assert a["nr"] == b["nr"]
    for k in ("pl", "sl", "radii"):
        np.testing.assert_array_equal(a[k], b[k])
    for x, y in zip(a["outs"] + a["grads"], b["outs"] + b["grads"]):
        assert rel_err(y, x) <= TOL


def test_fusion_plan_is_per_call_site_and_grad_mode_aware(torch_mod):
    """model_pn.py alternates cycles on ONE sampler (model_pn.py:768-786: interior samples -> u, ux, uxx; boundary samples ->
    u only; Model.forward: no_grad).  Each kind of cycle keeps its own fusion plan (keyed by the number of samples), pieces
    fused under no_grad are not handed out to a grad-enabled request, and an in-place edit of an input between forward
    and backward raises like autograd's own check."""
    torch = torch_mod
    from pigs_b200 import GaussianSampler, _lib
    rec = RECIPES["synth2d"]()
    d = to_dev(rec)
    bc = d["samples"][:777].contiguous()
    for k in ("means", "values", "conics"):
        d[k].requires_grad_(True)
    s = GaussianSampler(False, fused=True)
    L = _lib.lib()
    counts = {"int": [], "bc": []}
    for cycle in range(3):
        s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
        n0 = L.pigs_launch_count()
        u, du, ddu = s.sample_gaussians(), s.sample_gaussians_derivative(), s.sample_gaussians_laplacian()
        counts["int"].append(L.pigs_launch_count() - n0)
        s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], bc)
        n0 = L.pigs_launch_count()
        ub = s.sample_gaussians()
        counts["bc"].append(L.pigs_launch_count() - n0)
        assert ub.shape == (777, 1)
    assert counts["int"][1] < counts["int"][0] and counts["int"][2] == counts["int"][1]  # interior cycles fuse from the 2nd on
    assert counts["bc"][0] == counts["bc"][1] == counts["bc"][2]                         # the value-only cycle never inherits plan 7
    (u.sum() + du.sum()).backward()  # ddu unused: its gradient is not materialised, the backward runs mask 3
    assert d["means"].grad is not None
    # a fused evaluation started under no_grad does not serve a grad-enabled request
    s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
    with torch.no_grad():
        u0 = s.sample_gaussians()
    du1 = s.sample_gaussians_derivative()
    assert u0.grad_fn is None and du1.grad_fn is not None
    # in-place edit between forward and backward
    s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
    u2 = s.sample_gaussians()
    with torch.no_grad():
        d["values"].mul_(1.0)
    with pytest.raises(RuntimeError, match="inplace"):
        u2.sum().backward()
