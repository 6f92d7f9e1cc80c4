"""The C ABI driven by a compiled, torch-free host program (tests/abi_harness.cpp: cudaMalloc + the calls a maintainer of
the reference's sample_points.cu would make) — preprocess, fused forward, fused backward — compared
with the CPU oracle.  CPU: the harness builds with plain g++ against include/pigs_b200.h and the header parses as C."""
import os
import subprocess

import numpy as np
import pytest

from oracle import pigs_oracle as po
from tests.fixtures import synthetic
from tests.helpers import TOL, rel_err, upstream

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "tests", "_bin", "abi_harness")


def build_harness():
    os.makedirs(os.path.dirname(BIN), exist_ok=True)
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    cmd = ["g++", "-O2", "-std=c++17", os.path.join(ROOT, "tests", "abi_harness.cpp"), "-I" + os.path.join(ROOT, "include"),
           "-I" + os.path.join(cuda, "include"), "-L" + os.path.join(ROOT, "pigs_b200"), "-lpigs_b200",
           "-L" + os.path.join(cuda, "lib64"), "-lcudart", "-Wl,-rpath,$ORIGIN/../../pigs_b200",
           "-Wl,-rpath," + os.path.join(cuda, "lib64"), "-o", BIN]
    subprocess.run(cmd, check=True, capture_output=True, text=True)
    return BIN


def test_harness_builds_with_gpp_and_header_is_plain_c(tmp_path):
    build_harness()
    assert os.access(BIN, os.X_OK)
    c = tmp_path / "hdr.c"
    c.write_text('#include "pigs_b200.h"\nint main(void) { return (int)sizeof(pigs_header) - 256; }\n')
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-fsyntax-only", "-I" + os.path.join(ROOT, "include"), str(c)],
                   check=True, capture_output=True, text=True)


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [dict(P=3000, N=20000), dict(P=700, N=5000, D=1, C=2),
                                   dict(P=1500, N=8000, edge_fraction=0.2, anisotropy=4.0)])
def test_compiled_caller_matches_oracle(tmp_path, shape):
    if not os.path.exists(BIN):
        build_harness()
    kw = dict(shape)
    rec = synthetic(kw.pop("P"), kw.pop("N"), seed=31, **kw)
    P, D = rec["means"].shape
    N, C = rec["samples"].shape[0], rec["values"].shape[1]
    T = D * (D + 1) // 2
    gs = [upstream(rec, fn) for fn in range(3)]
    fin, fout = tmp_path / "in.bin", tmp_path / "out.bin"
    with open(fin, "wb") as f:
        np.array([P, D, N, C], np.int64).tofile(f)
        for a in (rec["means"], rec["covariances"], rec["conics"], rec["values"], rec["samples"], *gs):
            np.ascontiguousarray(a, np.float32).tofile(f)
    r = subprocess.run([BIN, str(fin), str(fout)], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr + r.stdout
    st = po.preprocess(rec["means"], rec["covariances"], rec["samples"])
    with open(fout, "rb") as f:
        R, blocks = np.fromfile(f, np.int64, 2)
        assert (R, blocks) == (st["num_rendered"], st["blocks"])
        take = lambda dt, n: np.fromfile(f, dt, int(n))
        np.testing.assert_array_equal(take(np.float32, P), st["radii"])
        np.testing.assert_array_equal(take(np.uint32, 2 * (blocks + 1)).reshape(-1, 2)[:blocks], st["ranges"][:blocks])
        np.testing.assert_array_equal(take(np.uint32, 2 * (blocks + 1)).reshape(-1, 2)[:blocks], st["sample_ranges"][:blocks])
        np.testing.assert_array_equal(take(np.uint32, R), st["point_list"])
        np.testing.assert_array_equal(take(np.uint32, N), st["sample_list"])
        a = (rec["means"], rec["values"], rec["conics"], rec["samples"])
        for fn, n in enumerate((N * C, N * D * C, N * D * D * C)):
            want = po.forward(fn, st, *a)
            assert rel_err(take(np.float32, n).reshape(want.shape), want) <= TOL, ("forward", fn)
        want = [np.zeros(s, np.float64) for s in ((P, D), (P, C), (P, T))]
        for fn in range(3):
            for acc, part in zip(want, po.backward(fn, st, *a, gs[fn])):
                acc += part.reshape(acc.shape)
        for w, name in zip(want, ("dmeans", "dvalues", "dconics")):
            assert rel_err(take(np.float32, w.size).reshape(w.shape), w) <= TOL, name
        assert f.read() == b""
