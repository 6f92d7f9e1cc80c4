"""The evaluation calls hold no host synchronisation (include/pigs_b200.h), so a forward / backward pair over a fixed
binning can be captured into a CUDA graph and replayed — the three class kernels fork onto library-owned streams and join
back before the call returns (csrc/pack.cu ClassFork), which stream capture follows.  Held here for both staging modes:
the replayed graph reproduces the eager results (1e-5 against the oracle; the forward's atomic order is not fixed), and
new input VALUES in the captured tensors flow through a replay."""
import numpy as np
import pytest

from oracle import pigs_oracle as po
from tests.fixtures import synthetic
from tests.helpers import TOL, oracle_state, rel_err, to_dev, upstream

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("P,N,pairs", [(1500, 9000, 1 << 30), (4000, 30000, 0)], ids=["single_class", "three_class"])
def test_forward_backward_replay_from_a_cuda_graph(P, N, pairs):
    import torch
    from pigs_b200 import _C, _lib
    from tests.conftest import SMALL_PAIRS_DEFAULT
    L = _lib.lib()
    L.pigs_set_small_problem_pairs(pairs)
    try:
        rec = synthetic(P, N, seed=31)
        d = to_dev(rec)
        ups = [torch.from_numpy(upstream(rec, fn)).cuda() for fn in range(3)]
        nr, bb, sbb, rg, srg, _ = _C.preprocess_gaussians(d["means"], d["values"], d["covariances"], d["conics"], d["samples"], False)
        a = (d["means"], d["values"], d["conics"], d["samples"], nr)

        def step():
            outs = _C.sample_forward(7, *a, bb, sbb, rg, srg, False)
            grads = _C.sample_backward(7, *a, [ups[0], ups[1], ups[2], None], bb, sbb, rg, srg, False)
            return outs, grads

        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):  # warm-up on the capture stream: the binding's scratch buffers are per stream
            for _ in range(3):
                step()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        n0 = L.pigs_launch_count()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            outs, grads = step()
        captured = L.pigs_launch_count() - n0
        assert captured == (5 if pairs else 9), captured  # pack + eval (+ sample pack) per call; 3 evaluation kernels each without the small mode

        st = oracle_state(rec)
        A = (rec["means"], rec["values"], rec["conics"], rec["samples"])

        def check(values):
            want_g = [np.zeros_like(rec[k], dtype=np.float64) for k in ("means", "values", "conics")]
            for fn in range(3):
                want = po.forward(fn, st, A[0], values, A[2], A[3])
                assert rel_err(outs[fn].cpu().numpy(), want) <= TOL, ("forward", fn)
                for acc, part in zip(want_g, po.backward(fn, st, A[0], values, A[2], A[3], upstream(rec, fn))):
                    acc += part.reshape(acc.shape)
            for got, want in zip(grads, want_g):
                assert rel_err(got.cpu().numpy().reshape(want.shape), want) <= TOL

        for o in list(outs[:3]) + list(grads):
            o.fill_(float("nan"))  # a replay must rewrite everything it owns (the zero fills are part of the graph)
        g.replay()
        torch.cuda.synchronize()
        check(rec["values"])
        # new values in the SAME tensor (the binning does not depend on them): the replay picks them up
        new_values = (rec["values"] * -0.5 + 0.25).astype(np.float32)
        d["values"].copy_(torch.from_numpy(new_values).cuda())
        g.replay()
        g.replay()
        torch.cuda.synchronize()
        check(new_values)
    finally:
        L.pigs_set_small_problem_pairs(SMALL_PAIRS_DEFAULT)
