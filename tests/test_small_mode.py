"""Small problems run ONE evaluation kernel per call (include/pigs_b200.h pigs_set_small_problem_pairs): every
(tile, Gaussian) instance takes the GENERAL class — the reference's exact per-pair torus wrap and `power > 0` skip,
forward.cu:149-171 — instead of three class kernels on forked streams.  Held here: the launch counts, the results against
the oracle and against the three-class path, a copied binning buffer (unknown to the library: three launches, still
right), and the threshold."""
import numpy as np
import pytest

from oracle import pigs_oracle as po
from tests.conftest import SMALL_PAIRS_DEFAULT
from tests.fixtures import RECIPES, synthetic
from tests.helpers import TOL, oracle_state, rel_err, to_dev, upstream

pytestmark = pytest.mark.gpu


def _run(rec, pairs, copy_buffers=False):
    """preprocess + fused u / grad u / Hessian forward + backward through the C ABI binding; returns outputs, gradients
    and the library launches of (preprocess, forward, backward)."""
    import torch
    from pigs_b200 import _C, _lib
    L = _lib.lib()
    L.pigs_set_small_problem_pairs(pairs)
    try:
        d = to_dev(rec)
        n0 = L.pigs_launch_count()
        nr, bb, sbb, rg, srg, radii = _C.preprocess_gaussians(d["means"], d["values"], d["covariances"], d["conics"],
                                                             d["samples"], False)
        n1 = L.pigs_launch_count()
        if copy_buffers:
            bb, sbb = bb.clone(), sbb.clone()
        a = (d["means"], d["values"], d["conics"], d["samples"], nr)
        outs = _C.sample_forward(7, *a, bb, sbb, rg, srg, False)
        n2 = L.pigs_launch_count()
        ups = [torch.from_numpy(upstream(rec, fn)).cuda() for fn in range(3)]
        grads = _C.sample_backward(7, *a, [ups[0], ups[1], ups[2], None], bb, sbb, rg, srg, False)
        n3 = L.pigs_launch_count()
        torch.cuda.synchronize()
        return [o.cpu().numpy() for o in outs[:3]], [g.cpu().numpy() for g in grads], (n1 - n0, n2 - n1, n3 - n2)
    finally:
        L.pigs_set_small_problem_pairs(SMALL_PAIRS_DEFAULT)


def _oracle(rec):
    st = oracle_state(rec)
    a = (rec["means"], rec["values"], rec["conics"], rec["samples"])
    outs = [po.forward(fn, st, *a) for fn in range(3)]
    grads = [np.zeros_like(rec[k], dtype=np.float64) for k in ("means", "values", "conics")]
    for fn in range(3):
        for acc, part in zip(grads, po.backward(fn, st, *a, upstream(rec, fn))):
            acc += part.reshape(acc.shape)
    return outs, grads


@pytest.mark.parametrize("name", ["synth2d", "torus", "edge_aniso", "wide", "line1d", "synth1d", "scaled"])
def test_single_class_matches_oracle_and_three_class(name):
    rec = RECIPES[name]()
    want_o, want_g = _oracle(rec)
    one_o, one_g, one_n = _run(rec, 1 << 30)
    three_o, three_g, three_n = _run(rec, 0)
    # launches: forward = pack + evaluation kernel(s); backward = pack + sample pack + evaluation kernel(s); the
    # preprocess of a small problem is two single-CTA launches (count stage, binning stage) either side of the readback
    assert one_n[1] == 2 and one_n[2] == 3, one_n
    assert three_n[1] == 4 and three_n[2] == 5, three_n
    # — while P, N <= 16 384 and num_rendered <= 8 192 (csrc/binning.cu SMALL_*_MAX); beyond, that stage keeps its grid-wide chain
    st = oracle_state(rec)
    P, N = rec["means"].shape[0], rec["samples"].shape[0]
    small_count = P <= 16384 and N <= 16384
    small_bin = small_count and st["num_rendered"] <= 8192 and 2 * st["blocks"] + 4 <= 256
    assert one_n[0] == (1 if small_count else 3) + (1 if small_bin else 9), (one_n, small_count, small_bin)
    assert three_n[0] == 13, three_n
    for got, other, want in zip(one_o + one_g, three_o + three_g, want_o + want_g):
        assert rel_err(got.reshape(want.shape), want) <= TOL
        assert rel_err(other.reshape(want.shape), want) <= TOL
        assert rel_err(got, other) <= TOL


def test_copied_binning_buffer_takes_three_launches_and_stays_right():
    """The host-side note is keyed by the buffer's address: a clone is unknown -> all three class kernels are launched
    (two of them on empty lists) and the result is the same."""
    rec = RECIPES["synth2d"]()
    want_o, want_g = _oracle(rec)
    got_o, got_g, n = _run(rec, 1 << 30, copy_buffers=True)
    assert n[1] == 4 and n[2] == 5, n
    for got, want in zip(got_o + got_g, want_o + want_g):
        assert rel_err(got.reshape(want.shape), want) <= TOL


def test_threshold_is_on_estimated_pairs():
    """60 K instances x 200 K samples / 16 tiles is far above the default: three classes; a threshold above the estimate
    switches the same problem to one class, with the same results."""
    rec = synthetic(20000, 200000, seed=11)
    big_o, big_g, big_n = _run(rec, SMALL_PAIRS_DEFAULT)
    assert big_n[1] == 4 and big_n[2] == 5, big_n
    one_o, one_g, one_n = _run(rec, 1 << 40)
    assert one_n[1] == 2 and one_n[2] == 3, one_n
    assert one_n[0] == big_n[0] - 1, (one_n, big_n)  # beyond the single-CTA limits: the grid-wide chain minus the class scan
    for a, b in zip(big_o + big_g, one_o + one_g):
        assert rel_err(a, b) <= TOL


def test_buffer_refilled_by_a_large_problem_forgets_the_small_note():
    """Same torch allocation reused: the note of the small binning must not survive a three-class preprocess into the
    same address (the lookup also compares num_rendered / N / blocks)."""
    import torch
    from pigs_b200 import GaussianSampler
    small, large = RECIPES["synth2d"](), synthetic(20000, 200000, seed=12)
    s = GaussianSampler(False)
    for rec in (small, large, small, large):
        d = to_dev(rec)
        s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
        st = oracle_state(rec)
        if rec is small:
            want = po.forward(0, st, rec["means"], rec["values"], rec["conics"], rec["samples"])
            assert rel_err(s.sample_gaussians().cpu().numpy(), want) <= TOL
        else:  # a subset of rows on the same tile grid (the pattern of tests/test_gpu_fullsize.py)
            rows = np.arange(0, rec["samples"].shape[0], 997)
            sub = np.ascontiguousarray(rec["samples"][rows])
            st_sub = po.preprocess(rec["means"], rec["covariances"], sub, bounds=(st["min_b"], st["max_b"]))
            want = po.forward(0, st_sub, rec["means"], rec["values"], rec["conics"], sub)
            assert rel_err(s.sample_gaussians().cpu().numpy()[rows], want) <= TOL
        del d
        torch.cuda.empty_cache()
