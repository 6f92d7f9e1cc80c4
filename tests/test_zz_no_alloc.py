"""pigs_neighbors_ranges_static — the allocation-free variant of pigs_neighbors_ranges (no counterpart in the reference;
added for callers that cannot call cudaMalloc: graph capture, serving loops).  Called through the C ABI (ctypes) and
held BIT-EXACT (int64 prefix sums) to the existing entry point on two calls in a row per size, since every call shares
the one static scratch array and must clear what it reuses.  Collected last on purpose.
"""
import ctypes

import pytest

torch = pytest.importorskip("torch")

from pigs_b200 import _lib  # noqa: E402

MAX_P = 1 << 24  # PIGS_RANGES_STATIC_MAX_P (include/pigs_b200.h)


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _existing(L, counts):
    P = counts.numel()
    ranges = torch.full((P,), -7, dtype=torch.int64, device="cuda")
    total = torch.full((1,), -7, dtype=torch.int64, device="cuda")
    ws = torch.empty(max(int(L.pigs_neighbors_scan_workspace_bytes(P)), 16), dtype=torch.uint8, device="cuda")
    _lib.check(L.pigs_neighbors_ranges(P, _ptr(counts), _ptr(ranges), _ptr(total), _ptr(ws), _stream()),
               "pigs_neighbors_ranges")
    return ranges, total


def _static(L, counts, ranges, total):
    _lib.check(L.pigs_neighbors_ranges_static(counts.numel(), _ptr(counts), _ptr(ranges), _ptr(total), _stream()),
               "pigs_neighbors_ranges_static")


@pytest.mark.gpu
@pytest.mark.parametrize("P", [1, 4097, 300_000, 1_000_003, 5_000_000, MAX_P])
def test_static_ranges_equal_the_existing_entry_point_twice_in_a_row(P):
    L = _lib.lib()
    torch.cuda.set_device(0)
    g = torch.Generator(device="cuda").manual_seed(1234 + P)
    ranges = torch.empty(P, dtype=torch.int64, device="cuda")
    total = torch.empty(1, dtype=torch.int64, device="cuda")
    for call in range(2):  # second call: other values, on the scratch the first call left behind
        counts = torch.randint(0, 1600 + 900 * call, (P,), generator=g, dtype=torch.int64, device="cuda")
        want_r, want_t = _existing(L, counts)
        ranges.fill_(-1)
        total.fill_(-1)
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        torch_alloc0 = torch.cuda.memory_allocated()
        n0 = L.pigs_launch_count()
        _static(L, counts, ranges, total)
        torch.cuda.synchronize()
        assert L.pigs_launch_count() == n0 + 1
        if call == 1:  # the first call on a device resolves the symbol (the driver may load the module then)
            assert torch.cuda.mem_get_info()[0] == free0, "the static entry point changed the device's free memory"
            assert torch.cuda.memory_allocated() == torch_alloc0
        assert torch.equal(ranges, want_r), (P, call)
        assert torch.equal(total, want_t) and int(total) == int(want_r[-1])


@pytest.mark.gpu
def test_static_ranges_back_to_back_without_a_sync_between_calls():
    """Stream order alone serialises two calls on one stream; both results must be right."""
    L = _lib.lib()
    torch.cuda.set_device(0)
    g = torch.Generator(device="cuda").manual_seed(99)
    a = torch.randint(0, 3000, (2_000_001,), generator=g, dtype=torch.int64, device="cuda")
    b = torch.randint(0, 50, (700_000,), generator=g, dtype=torch.int64, device="cuda")
    ra, ta = torch.empty_like(a), torch.empty(1, dtype=torch.int64, device="cuda")
    rb, tb = torch.empty_like(b), torch.empty(1, dtype=torch.int64, device="cuda")
    _static(L, a, ra, ta)
    _static(L, b, rb, tb)
    torch.cuda.synchronize()
    for counts, r, t in ((a, ra, ta), (b, rb, tb)):
        want_r, want_t = _existing(L, counts)
        assert torch.equal(r, want_r) and torch.equal(t, want_t)


@pytest.mark.gpu
def test_static_ranges_rejects_sizes_beyond_its_maximum():
    L = _lib.lib()
    torch.cuda.set_device(0)
    x = torch.zeros(16, dtype=torch.int64, device="cuda")
    rc = L.pigs_neighbors_ranges_static(MAX_P + 1, _ptr(x), _ptr(x), _ptr(x), _stream())
    assert rc < 0
    msg = L.pigs_last_error().decode()
    assert "exceeds" in msg and str(MAX_P) in msg, msg
    assert L.pigs_neighbors_ranges_static(0, _ptr(x), _ptr(x), _ptr(x), _stream()) < 0
    assert L.pigs_neighbors_ranges_static(16, None, _ptr(x), _ptr(x), _stream()) < 0
    torch.cuda.synchronize()


def test_static_maximum_matches_the_header():
    import os
    import re
    hdr = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "pigs_b200.h")).read()
    assert int(re.search(r"#define PIGS_RANGES_STATIC_MAX_P (\d+)", hdr).group(1)) == MAX_P
