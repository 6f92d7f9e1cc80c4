"""debug=True has the reference's semantics (cuda_sampler/auxiliary.h:33-40 CHECK_CUDA): the device is synchronised and
checked after EVERY kernel launch of the library, so a faulting kernel is reported at its own launch site.  The fault is
provoked in a child process (an illegal address poisons the CUDA context for good)."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu

_CHILD = r"""
import sys, ctypes
sys.path.insert(0, %r)
import torch
from pigs_b200 import _C, _lib
from tests.fixtures import synthetic
rec = synthetic(3000, 20000, seed=5)
d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
nr, bb, sbb, rg, srg, radii = _C.preprocess_gaussians(d["means"], d["values"], d["covariances"], d["conics"], d["samples"], True)
assert _lib.lib().pigs_get_debug() == 1
L = _lib.lib()
N, P = d["samples"].shape[0], d["means"].shape[0]
out = torch.zeros(N, 1, device="cuda")
ws_bytes = L.pigs_eval_workspace_bytes(nr, N, rg.numel() // 8 - 1, 2, 1, 1)
ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
p = lambda t: ctypes.c_void_p(t.data_ptr())
bad = ctypes.c_void_p(0x10)  # `means` points nowhere: the first kernel that reads it faults
rc = L.pigs_sample_forward(1, P, 2, N, 1, rg.numel() // 8 - 1, nr, bad, p(d["values"]), p(d["conics"]), p(d["samples"]),
                           p(bb), p(sbb), p(rg), p(srg), p(out), None, None, None, p(ws), ws_bytes,
                           ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
print("RC", rc)
print("ERR", L.pigs_last_error().decode())
"""


def test_debug_flag_sets_library_mode_and_keeps_results():
    import torch
    from pigs_b200 import GaussianSampler, _lib
    from tests.fixtures import synthetic
    rec = synthetic(2000, 9000, seed=3)
    d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
    outs = []
    for debug in (False, True):
        s = GaussianSampler(debug, fused=False)
        s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
        assert _lib.lib().pigs_get_debug() == (1 if debug else 0)
        outs.append((s.num_rendered, s.ranges.clone(), s.sample_ranges.clone(), s.sample_gaussians_derivative()))
    assert outs[0][0] == outs[1][0] and torch.equal(outs[0][1], outs[1][1]) and torch.equal(outs[0][2], outs[1][2])
    a, b = outs[0][3].cpu().numpy(), outs[1][3].cpu().numpy()
    assert np.abs(a - b).max() <= 2e-6 * np.abs(a).max()
    GaussianSampler(False).preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
    assert _lib.lib().pigs_get_debug() == 0


@pytest.mark.timeout(300)
def test_debug_mode_names_the_faulting_launch():
    r = subprocess.run([sys.executable, "-c", _CHILD % ROOT], capture_output=True, text=True, timeout=240)
    text = r.stdout + r.stderr
    assert "RC " in text, text[-2000:]
    rc = int(text.split("RC ")[1].split()[0])
    assert rc > 0, text[-2000:]  # a cudaError_t, returned by the call that launched the faulting kernel
    err = text.split("ERR ")[1].splitlines()[0]
    assert "debug: synchronised after this launch" in err and ".cu:" in err, err
