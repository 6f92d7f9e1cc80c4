"""Deterministic forward (include/pigs_b200.h pigs_set_deterministic): run-to-run BIT equality of u / grad u / Hessian, like
the reference forward (one thread per sample, Gaussians added in ascending id, forward.cu:100-166), plus the equivalence of
the concurrent and the in-order launch of the three class kernels."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run(d, mask=7):
    import torch
    from pigs_b200 import _C
    nr, bb, sbb, rg, srg, _ = _C.preprocess_gaussians(d["means"], d["values"], d["covariances"], d["conics"], d["samples"], False)
    outs = _C.sample_forward(mask, d["means"], d["values"], d["conics"], d["samples"], nr, bb, sbb, rg, srg, False)
    torch.cuda.synchronize()
    return [o for o in outs if o is not None]


@pytest.mark.parametrize("kw", [dict(P=20000, N=200000), dict(P=4000, N=60000, edge_fraction=0.15, anisotropy=4.0)])
def test_deterministic_forward_is_bitwise_reproducible_and_matches_default(kw):
    import torch
    import pigs_b200
    from tests.fixtures import synthetic
    kw = dict(kw)
    rec = synthetic(kw.pop("P"), kw.pop("N"), seed=31, **kw)
    d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
    default = _run(d)
    pigs_b200.set_deterministic(True)
    try:
        assert pigs_b200.get_deterministic()
        a = _run(d)
        b = _run(d)
        c = _run(d)
    finally:
        pigs_b200.set_deterministic(False)
    for x, y, z, w in zip(a, b, c, default):
        assert torch.equal(x, y) and torch.equal(x, z)          # bit for bit, run to run
        e = float((x - w).abs().max() / w.abs().max())
        assert e <= 1e-5, e                                      # same values as the default mode up to summation order


def test_deterministic_mode_matches_oracle():
    import torch
    import pigs_b200
    from oracle import pigs_oracle as po
    from tests.fixtures import RECIPES
    from tests.helpers import TOL, oracle_state, rel_err
    pigs_b200.set_deterministic(True)
    try:
        for name in ("synth2d", "torus", "edge_aniso", "synth1d", "synth2d_c2"):
            rec = RECIPES[name]()
            st = oracle_state(rec)
            d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
            outs = _run(d)
            for fn, got in enumerate(outs):
                w = po.forward(fn, st, rec["means"], rec["values"], rec["conics"], rec["samples"])
                assert rel_err(got.cpu().numpy(), w) <= TOL, (name, fn)
    finally:
        pigs_b200.set_deterministic(False)


def test_concurrent_and_in_order_class_kernels_agree():
    import torch
    from pigs_b200 import _C, _lib
    from tests.fixtures import synthetic
    from tests.helpers import upstream
    rec = synthetic(6000, 50000, seed=33, edge_fraction=0.2)
    d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
    gs = [torch.from_numpy(upstream(rec, fn)).cuda() for fn in range(3)] + [None]
    res = []
    L = _lib.lib()
    try:
        for on in (1, 0):
            L.pigs_set_class_concurrency(on)
            nr, bb, sbb, rg, srg, _ = _C.preprocess_gaussians(d["means"], d["values"], d["covariances"], d["conics"], d["samples"], False)
            a = (d["means"], d["values"], d["conics"], d["samples"], nr)
            outs = _C.sample_forward(7, *a, bb, sbb, rg, srg, False)
            grads = _C.sample_backward(7, *a, gs, bb, sbb, rg, srg, False)
            torch.cuda.synchronize()
            res.append([o for o in outs if o is not None] + list(grads))
    finally:
        L.pigs_set_class_concurrency(1)
    for x, y in zip(*res):
        e = float((x - y).abs().max() / y.abs().max())
        assert e <= 4e-6, e
