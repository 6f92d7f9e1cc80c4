"""Memory checks of our own, because compute-sanitizer is CLOSED on this GPU pool (profiles/r03a_sanitizer_closed_on_this_pool.txt:
"runs under it have left GPUs needing a reset ... find a bad access with bounds checks and asserts of your own, small cases,
and a comparison with the CPU reference").  Every buffer the binding hands to the library (workspaces, opaque binning buffers,
outputs, gradients) is allocated between two canary bands and, when the binding asked for uninitialised memory, pre-filled
with a poison pattern (0xFF bytes = NaN as float, huge as integer):
  * a write outside any buffer shows up as a damaged canary                      (memcheck's job),
  * a read of workspace memory nobody wrote shows up as NaN / garbage in results (initcheck's job),
  * results must still match the CPU oracle on the small recipes.
Shared-memory hazards (racecheck's job) are covered by tests/test_determinism.py: the deterministic mode must be bitwise
reproducible run to run, which a __syncthreads-ordering bug in the double-buffered staging would break."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GUARD = 4096  # bytes either side; keeps 128-byte sub-array alignment


class _Guarded:
    """Proxy for the `torch` module as seen by pigs_b200._C: allocation functions return the inside of a canary sandwich."""

    def __init__(self, torch):
        self._t = torch
        self.blocks = []

    def __getattr__(self, k):
        return getattr(self._t, k)

    def _alloc(self, shape, dtype, device, fill):
        t = self._t
        shape = tuple(int(s) for s in (shape if isinstance(shape, (tuple, list, t.Size)) else (shape,)))
        n = int(np.prod(shape)) if len(shape) else 1
        item = t.empty(0, dtype=dtype).element_size()
        raw = t.full((2 * GUARD + n * item,), 0xA5, dtype=t.uint8, device=device)
        inner = raw[GUARD:GUARD + n * item]
        if fill is None:
            inner.fill_(0xFF)  # poison: NaN as fp32, -1 as int
        else:
            inner.zero_()
        self.blocks.append(raw)
        out = inner.view(dtype).view(shape)
        if fill not in (None, 0):
            out.fill_(fill)
        return out

    def empty(self, *shape, dtype=None, device=None, **kw):
        if device is None or str(device) == "cpu" or kw.get("pin_memory"):
            return self._t.empty(*shape, dtype=dtype, device=device, **kw)
        shape = shape[0] if len(shape) == 1 and not isinstance(shape[0], int) else shape
        return self._alloc(shape, dtype or self._t.float32, device, None)

    def zeros(self, *shape, dtype=None, device=None, **kw):
        if device is None or str(device) == "cpu":
            return self._t.zeros(*shape, dtype=dtype, device=device, **kw)
        shape = shape[0] if len(shape) == 1 and not isinstance(shape[0], int) else shape
        return self._alloc(shape, dtype or self._t.float32, device, 0)

    def full(self, shape, value, dtype=None, device=None, **kw):
        if device is None or str(device) == "cpu":
            return self._t.full(shape, value, dtype=dtype, device=device, **kw)
        return self._alloc(shape, dtype or self._t.float32, device, value)

    def zeros_like(self, x, **kw):
        return self._alloc(tuple(x.shape), x.dtype, x.device, 0) if x.is_cuda else self._t.zeros_like(x, **kw)

    def empty_like(self, x, **kw):
        return self._alloc(tuple(x.shape), x.dtype, x.device, None) if x.is_cuda else self._t.empty_like(x, **kw)

    def check(self):
        self._t.cuda.synchronize()
        for raw in self.blocks:
            lo, hi = raw[:GUARD], raw[-GUARD:]
            assert bool((lo == 0xA5).all()) and bool((hi == 0xA5).all()), "canary damaged: out-of-bounds write"
        return len(self.blocks)


@pytest.fixture
def guarded(monkeypatch):
    import torch
    from pigs_b200 import _C
    g = _Guarded(torch)
    monkeypatch.setattr(_C, "torch", g)
    _C.use_binding("ctypes")  # the binding whose allocations go through the guarded proxy
    yield g
    if _C._ext is not None:
        _C.use_binding("compiled")


@pytest.mark.parametrize("name", ["synth2d", "torus", "wide", "edge_aniso", "synth1d", "synth2d_c2", "line1d"])
def test_sampling_path_stays_inside_its_buffers(name, guarded):
    import torch
    from oracle import pigs_oracle as po
    from pigs_b200 import _C
    from tests.fixtures import RECIPES
    from tests.helpers import TOL, oracle_state, ranges_np, rel_err, upstream
    rec = RECIPES[name]()
    st = oracle_state(rec)
    d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
    nr, bb, sbb, rg, srg, radii = _C.preprocess_gaussians(d["means"], d["values"], d["covariances"], d["conics"], d["samples"], False)
    assert nr == st["num_rendered"]
    np.testing.assert_array_equal(ranges_np(rg, st["blocks"]), st["ranges"])
    np.testing.assert_array_equal(ranges_np(srg, st["blocks"]), st["sample_ranges"])
    a = (d["means"], d["values"], d["conics"], d["samples"], nr)
    for mask in (15, 1 | 2 | 16):
        outs = _C.sample_forward(mask, *a, bb, sbb, rg, srg, False)
        for fn in range(4):
            if mask & (1 << fn) and not (fn == 2 and mask & 16):
                w = po.forward(fn, st, rec["means"], rec["values"], rec["conics"], rec["samples"])
                assert rel_err(outs[fn].cpu().numpy(), w) <= TOL, (name, mask, fn)
    gs = [torch.from_numpy(upstream(rec, fn)).cuda() for fn in range(4)]
    dm, dv, dc = _C.sample_backward(15, *a, gs, bb, sbb, rg, srg, False)
    want = [0, 0, 0]
    for fn in range(4):
        for k, part in enumerate(po.backward(fn, st, rec["means"], rec["values"], rec["conics"], rec["samples"], upstream(rec, fn))):
            want[k] = want[k] + part.astype(np.float64)
    for got, w in zip((dm, dv, dc), want):
        assert rel_err(got.cpu().numpy().reshape(w.shape), w) <= TOL, name
    assert guarded.check() > 8


@pytest.mark.parametrize("name", ["agg_model", "agg_model_1d", "agg_grid5"])
def test_aggregation_path_stays_inside_its_buffers(name, guarded):
    import torch
    from oracle import pigs_oracle as po
    from pigs_b200 import _C
    from tests.fixtures import AGG_RECIPES
    from tests.helpers import AGG_KEYS, TOL, agg_upstream, oracle_state, rel_err
    rec = AGG_RECIPES[name]()
    st = oracle_state(rec)
    d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
    radii = torch.from_numpy(st["radii"]).cuda()
    idx, rng, dists, dens, inv = _C.preprocess_aggregate(d["means"], d["conics"], radii, False)
    oi, orng, od, oden, oinv = po.neighbors(rec["means"], rec["conics"], st["radii"])
    np.testing.assert_array_equal(idx.cpu().numpy(), oi)
    np.testing.assert_array_equal(rng.cpu().numpy(), orng)
    a = [d[k] for k in AGG_KEYS]
    w, e, f, out = _C.aggregate_neighbors(*a, idx, rng, dists, dens, inv, False)
    ow, oe, of, oout = po.aggregate_forward(*[rec[k] for k in AGG_KEYS], oi, orng, od, oden, oinv)
    assert rel_err(out.cpu().numpy(), oout) <= TOL
    up = agg_upstream(rec)
    grads = _C.aggregate_neighbors_backward(*a, idx, rng, dists, dens, w, e, f, inv, torch.from_numpy(up).cuda(), False)
    want = po.aggregate_backward(*[rec[k] for k in AGG_KEYS], oi, orng, od, oden, ow, oe, of, oinv, up)
    for key, g, wv in zip(AGG_KEYS, grads, want):
        assert rel_err(g.cpu().numpy(), wv) <= TOL, key
    assert guarded.check() > 8
