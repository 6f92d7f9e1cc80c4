import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "slow: long comparison against the live reference extension (PIGS_RUN_SLOW=1)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_cuda = torch.cuda.is_available()
    except Exception:
        has_cuda = False
    if has_cuda:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


# Small problems run ONE evaluation kernel (every instance in the GENERAL class, pigs_set_small_problem_pairs), single-CTA
# binning kernels and a brute-force neighbour search (pigs_set_neighbor_small); the small recipes of these modules would
# otherwise stop exercising the FAST / SHIFT kernels, the grid-wide binning chain and the grid search, so they run in both
# modes ("three_class" = every small-problem path off).
_BOTH_CLASS_MODES = {"test_gpu_parity.py", "test_gpu_vs_reference.py", "test_culled.py", "test_gpu_aggregate.py"}
SMALL_PAIRS_DEFAULT = 20_000_000
NB_SMALL_P_DEFAULT = 4096


def pytest_generate_tests(metafunc):
    if os.path.basename(metafunc.module.__file__) in _BOTH_CLASS_MODES and "_class_mode" in metafunc.fixturenames:
        metafunc.parametrize("_class_mode", ["single_class", "three_class"], indirect=True)


@pytest.fixture(autouse=True)
def _class_mode(request):
    mode = getattr(request, "param", None)
    if mode != "three_class":
        yield mode
        return
    from pigs_b200 import _lib
    L = _lib.lib()
    L.pigs_set_small_problem_pairs(0)
    L.pigs_set_neighbor_small(0)
    try:
        yield mode
    finally:
        L.pigs_set_small_problem_pairs(SMALL_PAIRS_DEFAULT)
        L.pigs_set_neighbor_small(NB_SMALL_P_DEFAULT)
