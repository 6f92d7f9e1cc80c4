"""The compiled torch binding (pigs_b200/_Cext.so, csrc/torch_ext.cpp — the counterpart of the reference's pybind module,
ext.cpp:19-32) against the ctypes binding over the same C ABI: bit-identical bins, equal outputs and gradients."""
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_NAMES = ["preprocess_gaussians", "sample_gaussians", "sample_gaussians_backward", "sample_gaussians_derivative",
             "sample_gaussians_derivative_backward", "sample_gaussians_laplacian", "sample_gaussians_laplacian_backward",
             "sample_gaussians_third_derivative", "sample_gaussians_third_derivative_backward", "preprocess_aggregate",
             "aggregate_neighbors", "aggregate_neighbors_backward"]  # ext.cpp:20-31


def test_compiled_module_is_built_and_exports_the_reference_entry_points():
    from pigs_b200 import _C
    assert os.path.exists(os.path.join(ROOT, "pigs_b200", "_Cext.so")), "build() compiles the torch binding"
    assert _C._ext is not None and _C.BINDING == "compiled"
    for n in REF_NAMES + ["sample_forward", "sample_backward"]:
        assert callable(getattr(_C._ext, n)), n
        assert callable(getattr(_C, n + "_ctypes")), n
    import diff_gaussian_sampling
    assert diff_gaussian_sampling._C is _C  # the reference's `from . import _C` surface


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["synth2d", "torus", "synth1d", "synth2d_c2"])
def test_compiled_and_ctypes_bindings_agree(name):
    import torch
    from pigs_b200 import _C
    from tests.fixtures import RECIPES
    from tests.helpers import upstream
    rec = RECIPES[name]()
    d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
    gs = [torch.from_numpy(upstream(rec, fn)).cuda() for fn in range(4)]
    res = {}
    try:
        for b in ("compiled", "ctypes"):
            _C.use_binding(b)
            nr, bb, sbb, rg, srg, radii = _C.preprocess_gaussians(d["means"], d["values"], d["covariances"], d["conics"],
                                                                 d["samples"], False)
            N = d["samples"].shape[0]
            pl, pt, sl = _C.point_lists(nr, bb, sbb, N)
            a = (d["means"], d["values"], d["conics"], d["samples"], nr)
            outs = _C.sample_forward(15, *a, bb, sbb, rg, srg, False)
            single = _C.sample_gaussians_derivative(*a, bb, sbb, rg, srg, False)
            grads = _C.sample_backward(15, *a, gs, bb, sbb, rg, srg, False)
            sg = _C.sample_gaussians_laplacian_backward(*a, gs[2], bb, sbb, rg, srg, False)
            assert _C.packed_base(grads[0]).numel() == sum(g.numel() for g in grads)
            res[b] = dict(nr=nr, rg=rg.clone(), srg=srg.clone(), radii=radii.clone(), pl=pl.clone(), pt=pt.clone(), sl=sl.clone(),
                          outs=[o.clone() for o in outs], single=single.clone(), grads=[g.clone() for g in grads],
                          sg=[g.clone() for g in sg])
    finally:
        _C.use_binding("compiled")
    c, t = res["compiled"], res["ctypes"]
    assert c["nr"] == t["nr"]
    for k in ("rg", "srg", "radii", "pl", "pt", "sl"):
        assert torch.equal(c[k], t[k]), k                       # bins: bit for bit
    close = lambda x, y: float((x - y).abs().max()) <= 4e-6 * float(y.abs().max() + 1e-30)
    for x, y in zip(c["outs"] + [c["single"]] + c["grads"] + c["sg"], t["outs"] + [t["single"]] + t["grads"] + t["sg"]):
        assert x.shape == y.shape and close(x, y)               # values: up to the order of the atomic adds


@pytest.mark.gpu
def test_compiled_aggregation_matches_ctypes_and_errors_like_the_reference():
    import torch
    from pigs_b200 import _C
    from tests.fixtures import AGG_RECIPES
    from tests.helpers import AGG_KEYS, agg_upstream, oracle_state
    rec = AGG_RECIPES["agg_model"]()
    st = oracle_state(rec)
    d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
    radii = torch.from_numpy(st["radii"]).cuda()
    a = [d[k] for k in AGG_KEYS]
    up = torch.from_numpy(agg_upstream(rec)).cuda()
    res = {}
    try:
        for b in ("compiled", "ctypes"):
            _C.use_binding(b)
            csr = _C.preprocess_aggregate(d["means"], d["conics"], radii, False)
            half = _C.preprocess_aggregate(d["means"], d["conics"], radii, False, rows=(10, 200))
            w, e, f, out = _C.aggregate_neighbors(*a, *csr, False)
            grads = _C.aggregate_neighbors_backward(*a, csr[0], csr[1], csr[2], csr[3], w, e, f, csr[4], up, False)
            res[b] = (csr, half, out, grads)
    finally:
        _C.use_binding("compiled")
    for x, y in zip(res["compiled"][0] + res["compiled"][1], res["ctypes"][0] + res["ctypes"][1]):
        assert torch.equal(x, y)                                # neighbour lists: bit for bit
    assert torch.equal(res["compiled"][2], res["ctypes"][2])
    for x, y in zip(res["compiled"][3], res["ctypes"][3]):
        assert float((x - y).abs().max()) <= 1e-5 * float(y.abs().max() + 1e-30)
    with pytest.raises(RuntimeError):   # the reference's .data<float>() throws on a double tensor
        _C.preprocess_gaussians(d["means"].double(), d["values"], d["covariances"], d["conics"], d["samples"], False)
    with pytest.raises(RuntimeError):
        _C.preprocess_gaussians(d["means"].cpu(), d["values"], d["covariances"], d["conics"], d["samples"], False)
