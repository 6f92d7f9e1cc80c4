"""The C-ABI library loads without a GPU and exports every symbol include/pigs_b200.h declares (no compute calls)."""
import ctypes
import os
import re

from pigs_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "pigs_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(pigs_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported_and_bound():
    names = _declared()
    assert len(names) >= 25
    so = os.path.join(ROOT, "pigs_b200", "libpigs_b200.so")
    if not os.path.exists(so):
        _lib.build()
    L = ctypes.CDLL(so)
    for n in names:
        assert hasattr(L, n), "libpigs_b200.so does not export %s" % n
        assert n in _lib.SIGNATURES, "pigs_b200/_lib.py has no binding for %s" % n
    assert set(_lib.SIGNATURES) <= set(names), set(_lib.SIGNATURES) - set(names)


def test_host_only_queries():
    L = _lib.lib()
    assert L.pigs_version() >= 100
    assert L.pigs_geom_workspace_bytes(1000) > 4000
    assert L.pigs_binning_bytes(5000, 16) >= 2 * 4 * 5000 + 256
    assert L.pigs_sample_binning_bytes(5000, 16, 2) >= 2 * 4 * 5000 + 256
    assert L.pigs_eval_workspace_bytes(5000, 3000, 16, 2, 1, 7) > 0
    assert ctypes.sizeof(_lib.PigsHeader) == 256


def test_no_oracle_import_in_product():
    """The product never routes through the oracle (or any CPU fallback)."""
    pkg = os.path.join(ROOT, "pigs_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "pigs_oracle" not in txt and "from oracle" not in txt and "import oracle" not in txt, f
