"""ctypes loader for libpigs_b200.so (the C ABI in include/pigs_b200.h).

There is no CPU fallback and no alternative backend: if the shared library is missing or a CUDA
call fails, a RuntimeError is raised.
"""
import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
# PIGS_B200_LIB: load another build of the same library (tuning experiments: tools/*.py time variant builds)
_SO = os.environ.get("PIGS_B200_LIB") or os.path.join(_HERE, "libpigs_b200.so")
_lib = None

c_void_p, c_int, c_int64, c_size_t, c_uint = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_size_t, ctypes.c_uint


class PigsHeader(ctypes.Structure):
    """Mirror of `struct pigs_header` (include/pigs_b200.h)."""
    _fields_ = [
        ("magic", ctypes.c_uint32),
        ("D", ctypes.c_int32),
        ("grid", ctypes.c_int32 * 2),
        ("blocks", ctypes.c_int32),
        ("nslots", ctypes.c_int32),
        ("min_b", ctypes.c_float * 2),
        ("max_b", ctypes.c_float * 2),
        ("P", ctypes.c_int64),
        ("N", ctypes.c_int64),
        ("num_rendered", ctypes.c_int64),
        ("tile_width", ctypes.c_float),
        ("cull_sigmas", ctypes.c_float),
        ("subdiv", ctypes.c_int32),
        ("ref_grid", ctypes.c_int32 * 2),
        ("reserved", ctypes.c_uint32 * 43),
    ]


HEADER_BYTES = ctypes.sizeof(PigsHeader)
assert HEADER_BYTES == 256, HEADER_BYTES

# name -> (restype, argtypes); every symbol include/pigs_b200.h declares
SIGNATURES = {
    "pigs_version": (c_int, []),
    "pigs_last_error": (ctypes.c_char_p, []),
    "pigs_set_grid_recip_mode": (None, [c_int]),
    "pigs_sample_bounds": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_void_p]),
    "pigs_geom_workspace_bytes": (c_size_t, [c_int64]),
    "pigs_preprocess_count": (c_int, [c_int64, c_int, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "pigs_preprocess_count_culled": (c_int, [c_int64, c_int, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                             c_int, ctypes.c_float, c_void_p]),
    "pigs_preprocess_count_fused": (c_int, [c_int64, c_int, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p,
                                            c_void_p, c_int, ctypes.c_float, c_void_p]),
    "pigs_binning_bytes": (c_size_t, [c_int64, c_int]),
    "pigs_sample_binning_bytes": (c_size_t, [c_int64, c_int, c_int]),
    "pigs_sort_workspace_bytes": (c_size_t, [c_int64, c_int64]),
    "pigs_preprocess_bin": (c_int, [c_int64, c_int, c_int64, c_int, c_int64] + [c_void_p] * 11),
    "pigs_point_list": (c_void_p, [c_void_p]),
    "pigs_point_tiles": (c_void_p, [c_void_p, c_int64]),
    "pigs_sample_list": (c_void_p, [c_void_p]),
    "pigs_eval_workspace_bytes": (c_size_t, [c_int64, c_int64, c_int, c_int, c_int, c_uint]),
    "pigs_sample_forward": (c_int, [c_uint, c_int64, c_int, c_int64, c_int, c_int, c_int64] + [c_void_p] * 13 + [c_size_t, c_void_p]),
    "pigs_sample_backward": (c_int, [c_uint, c_int64, c_int, c_int64, c_int, c_int, c_int64] + [c_void_p] * 16 + [c_size_t, c_void_p]),
    "pigs_launch_count": (c_int64, []),
    "pigs_set_debug": (None, [c_int]),
    "pigs_get_debug": (c_int, []),
    "pigs_set_deterministic": (None, [c_int]),
    "pigs_get_deterministic": (c_int, []),
    "pigs_set_class_concurrency": (None, [c_int]),
    "pigs_set_small_problem_pairs": (None, [c_int64]),
    "pigs_set_stage_timing": (None, [c_int]),
    "pigs_stage_timing_read": (c_int, [c_void_p, c_void_p, c_int]),
    "pigs_set_kernel_timing": (None, [c_int]),
    "pigs_kernel_timing_read": (c_int, [c_int, c_void_p, c_void_p]),
    "pigs_neighbor_workspace_bytes": (c_size_t, [c_int64, c_int]),
    "pigs_neighbors_count": (c_int, [c_int64, c_int] + [c_void_p] * 5),
    "pigs_neighbors_fill": (c_int, [c_int64, c_int] + [c_void_p] * 10),
    "pigs_neighbors_scan_workspace_bytes": (c_size_t, [c_int64]),
    "pigs_neighbors_ranges": (c_int, [c_int64] + [c_void_p] * 5),
    "pigs_neighbors_ranges_static": (c_int, [c_int64] + [c_void_p] * 4),
    "pigs_neighbors_count_rows": (c_int, [c_int64, c_int, c_void_p, c_void_p, c_int64, c_int64] + [c_void_p] * 3),
    "pigs_neighbors_fill_rows": (c_int, [c_int64, c_int, c_void_p, c_void_p, c_void_p, c_int64, c_int64] + [c_void_p] * 7),
    "pigs_aggregate_forward": (c_int, [c_int64, c_int, c_int, c_int, c_int] + [c_void_p] * 16),
    "pigs_aggregate_forward_rows": (c_int, [c_int64, c_int, c_int, c_int, c_int, c_int64, c_int64] + [c_void_p] * 16),
    "pigs_aggregate_backward_rows": (c_int, [c_int64, c_int, c_int, c_int, c_int, c_int64, c_int64, c_int64] + [c_void_p] * 23),
    "pigs_aggregate_forward_f64": (c_int, [c_int64, c_int, c_int, c_int, c_int] + [c_void_p] * 16),
    "pigs_aggregate_backward_f64": (c_int, [c_int64, c_int, c_int, c_int, c_int] + [c_void_p] * 22),
    "pigs_set_aggregate_specialised": (None, [c_int]),
    "pigs_set_neighbor_search": (None, [c_int, ctypes.c_float]),
    "pigs_set_neighbor_small": (None, [c_int64]),
    "pigs_aggregate_backward_workspace_bytes": (c_size_t, [c_int64, c_int, c_int, c_int]),
    "pigs_aggregate_backward": (c_int, [c_int64, c_int, c_int, c_int, c_int, c_int64] + [c_void_p] * 23),
    "pigs_build_covariances_forward": (c_int, [c_int64, c_int] + [c_void_p] * 5),
    "pigs_build_covariances_backward": (c_int, [c_int64, c_int] + [c_void_p] * 7),
    "pigs_pde_residual_workspace_bytes": (c_size_t, [c_int64, c_int]),
    "pigs_pde_residual_forward": (c_int, [c_int, c_int64, c_int, c_int, ctypes.c_float, ctypes.c_float] + [c_void_p] * 8),
    "pigs_pde_residual_backward": (c_int, [c_int, c_int64, c_int, c_int, ctypes.c_float, ctypes.c_float] + [c_void_p] * 10),
    "pigs_pde_ns_residual_forward": (c_int, [c_int64, ctypes.c_float, ctypes.c_float] + [c_void_p] * 10),
    "pigs_pde_ns_residual_backward": (c_int, [c_int64, ctypes.c_float, ctypes.c_float] + [c_void_p] * 12),
    "pigs_pde_residual_timelevels_forward": (c_int, [c_int, c_int64, c_int, c_int, ctypes.c_float, ctypes.c_float, c_void_p,
                                                     ctypes.c_float] + [c_void_p] * 10),
    "pigs_pde_residual_timelevels_backward": (c_int, [c_int, c_int64, c_int, c_int, ctypes.c_float, ctypes.c_float, c_void_p,
                                                      ctypes.c_float] + [c_void_p] * 14),
    "pigs_pde_ns_residual_timelevels_forward": (c_int, [c_int64, ctypes.c_float, ctypes.c_float, c_void_p, ctypes.c_float]
                                                + [c_void_p] * 15),
    "pigs_pde_ns_residual_timelevels_backward": (c_int, [c_int64, ctypes.c_float, ctypes.c_float, c_void_p, ctypes.c_float]
                                                 + [c_void_p] * 22),
    "pigs_set_dense_slots": (None, [c_int, c_int]),
    "pigs_dense_forward": (c_int, [ctypes.c_uint, c_int64, c_int, c_int, c_int64] + [c_void_p] * 8),
    "pigs_dense_backward": (c_int, [ctypes.c_uint, c_int64, c_int, c_int, c_int64] + [c_void_p] * 11),
    "pigs_vorticity_forward": (c_int, [c_int64, c_void_p, c_void_p, c_void_p]),
    "pigs_vorticity_backward": (c_int, [c_int64, c_void_p, c_void_p, c_void_p]),
    "pigs_compact_workspace_bytes": (c_size_t, [c_int64]),
    "pigs_compact_count": (c_int, [c_int64, c_void_p, c_void_p, c_void_p, c_void_p]),
    "pigs_compact_fill": (c_int, [c_int64, c_void_p, c_void_p, c_int, c_void_p, c_int64, c_void_p, c_void_p]),
    "pigs_gather_rows": (c_int, [c_int64, c_int, c_int64, c_void_p, c_void_p, c_void_p, ctypes.c_float, c_void_p]),
    "pigs_split_means": (c_int, [c_int64, c_int, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "pigs_microbench_ffma": (c_int64, [c_int, c_int, c_void_p, c_void_p]),
    "pigs_microbench_ex2": (c_int64, [c_int, c_int, c_void_p, c_void_p]),
    "pigs_microbench_ffma2": (c_int64, [c_int, c_int, c_int, c_void_p, c_void_p]),
}


def build(verbose=False):
    """Compile pigs_b200/csrc/*.cu for sm_100a into pigs_b200/libpigs_b200.so (nvcc, in-tree)."""
    cmd = ["make", "-C", os.path.join(_HERE, "csrc"), "-j8"]
    if not verbose:
        cmd.append("-s")
    subprocess.run(cmd, check=True)


def build_ext(verbose=False):
    """Compile the torch binding pigs_b200/_Cext.so (csrc/torch_ext.cpp, host compiler only) over the library."""
    subprocess.run(["bash", os.path.join(_HERE, "csrc", "build_ext.sh")], check=True,
                   stdout=None if verbose else subprocess.DEVNULL)


def lib():
    """The loaded library; raises RuntimeError when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            raise RuntimeError(
                f"{_SO} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a). pigs_b200 has no CPU or PyTorch fallback.")
        L = ctypes.CDLL(_SO)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError if the .so does not export a declared symbol
            fn.restype = res
            fn.argtypes = args
        if os.environ.get("PIGS_DETERMINISTIC", "0") not in ("", "0"):
            L.pigs_set_deterministic(1)  # bitwise reproducible forward (include/pigs_b200.h)
        _lib = L
    return _lib


def check(rc, what):
    if rc != 0:
        msg = lib().pigs_last_error()
        raise RuntimeError(f"pigs_b200: {what} failed (status {rc}): {msg.decode() if msg else ''}")
