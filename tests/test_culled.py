"""Opt-in culled mode: every reference tile cut k x k and a `sigmas`-standard-deviation box
test per fine tile.  It evaluates a SUBSET of the reference's pair set, so it is held to two separate statements:
(1) it is the same pipeline on different bins — bins bit-exact and values / gradients within 1e-5 of the oracle run
with the same two parameters; (2) its deviation from the exact mode is bounded by what the dropped pairs can
contribute (measured at full size by tools/bench_culled.py)."""
import numpy as np
import pytest

from oracle import pigs_oracle as po
from tests.fixtures import synthetic
from tests.helpers import TOL, ranges_np, rel_err, upstream

CULLS = [(6.0, 4), (5.0, 8), (6.0, 1), (4.0, 3)]


def _rec(**kw):
    return synthetic(6000, 20000, seed=21, **kw)


@pytest.mark.parametrize("cull", CULLS)
def test_oracle_culled_mode_is_a_subset_with_small_deviation(cull):
    rec = _rec()
    a = (rec["means"], rec["values"], rec["conics"], rec["samples"])
    st0 = po.preprocess(rec["means"], rec["covariances"], rec["samples"])
    st1 = po.preprocess(rec["means"], rec["covariances"], rec["samples"], cull=cull)
    if cull[1] == 1:
        assert po.pair_count(st1) == po.pair_count(st0)  # no subdivision: the reference pair set itself
        np.testing.assert_array_equal(st1["point_list"], st0["point_list"])
    else:
        assert po.pair_count(st1) < po.pair_count(st0)
    np.testing.assert_array_equal(st1["radii"], st0["radii"])  # radii stay the reference's
    # a dropped pair is worth < exp(-sigmas^2/2) * |v| * (polynomial in a): 4 sigma 3.4e-4, 5 sigma 3.7e-6, 6 sigma 1.5e-8
    bound = {4.0: 2e-2, 5.0: 2e-4, 6.0: 1e-5}[cull[0]]
    for fn in range(3):
        exact = po.forward(fn, st0, *a)
        culled = po.forward(fn, st1, *a)
        assert rel_err(culled, exact) <= bound, (fn, rel_err(culled, exact))


def test_oracle_default_is_untouched_after_a_culled_call():
    rec = _rec()
    st0 = po.preprocess(rec["means"], rec["covariances"], rec["samples"])
    po.preprocess(rec["means"], rec["covariances"], rec["samples"], cull=(5.0, 4))
    st2 = po.preprocess(rec["means"], rec["covariances"], rec["samples"])
    np.testing.assert_array_equal(st0["radii"], st2["radii"])
    np.testing.assert_array_equal(st0["point_list"], st2["point_list"])


@pytest.mark.gpu
@pytest.mark.parametrize("cull", CULLS)
@pytest.mark.parametrize("kw", [dict(), dict(edge_fraction=0.2, anisotropy=4.0)])
def test_cuda_culled_matches_oracle_culled(cull, kw):
    import torch

    from diff_gaussian_sampling import GaussianSampler
    from pigs_b200 import _C
    rec = _rec(**kw)
    a = (rec["means"], rec["values"], rec["conics"], rec["samples"])
    st = po.preprocess(rec["means"], rec["covariances"], rec["samples"], cull=cull)
    d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
    for k in ("means", "values", "conics"):
        d[k].requires_grad_(True)
    s = GaussianSampler(False, cull=cull)
    s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
    assert s.num_rendered == st["num_rendered"]
    np.testing.assert_array_equal(s.radii.cpu().numpy(), st["radii"])
    np.testing.assert_array_equal(ranges_np(s.ranges, st["blocks"]), st["ranges"])
    np.testing.assert_array_equal(ranges_np(s.sample_ranges, st["blocks"]), st["sample_ranges"])
    pl, _, sl = _C.point_lists(s.num_rendered, s.binning_buffer, s.sample_binning_buffer, rec["samples"].shape[0])
    np.testing.assert_array_equal(pl.cpu().numpy().view(np.uint32), st["point_list"])
    np.testing.assert_array_equal(sl.cpu().numpy().view(np.uint32), st["sample_list"])
    assert s.pair_count() == po.pair_count(st)
    outs = s.sample_fused(7)
    gs = [upstream(rec, fn) for fn in range(3)]
    want = [np.zeros(rec[k].shape, np.float64) for k in ("means", "values", "conics")]
    for fn, got in enumerate(outs):
        assert rel_err(got.detach().cpu().numpy(), po.forward(fn, st, *a)) <= TOL, fn
        for acc, part in zip(want, po.backward(fn, st, *a, gs[fn])):
            acc += part.reshape(acc.shape)
    sum((o * torch.from_numpy(g).cuda()).sum() for o, g in zip(outs, gs)).backward()
    for k, w in zip(("means", "values", "conics"), want):
        assert rel_err(d[k].grad.cpu().numpy().reshape(w.shape), w) <= TOL, k


@pytest.mark.gpu
def test_cuda_culled_auto_width_and_deviation_from_exact():
    import torch

    from diff_gaussian_sampling import GaussianSampler
    rec = synthetic(40000, 300000, seed=22)
    d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
    exact = GaussianSampler(False)
    exact.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
    culled = GaussianSampler(False, cull=6.0)
    culled.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
    assert culled.pair_count() < 0.8 * exact.pair_count()
    for a, b in zip(culled.sample_fused(7), exact.sample_fused(7)):
        assert rel_err(a.cpu().numpy(), b.cpu().numpy()) <= TOL
