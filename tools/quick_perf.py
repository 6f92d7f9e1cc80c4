"""Dev tool: microbenchmarks + a quick timing of the fused forward/backward (run on the GPU box)."""
import ctypes
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pigs_b200 import _lib  # noqa: E402
for _i, _a in enumerate(sys.argv):  # --lib <path>: time a debug build of the library (issue-rate experiments)
    if _a == "--lib":
        _lib._SO = os.path.abspath(sys.argv[_i + 1])
from pigs_b200 import GaussianSampler  # noqa: E402
from tests.fixtures import synthetic  # noqa: E402


def ev_time(fn, warm=2, it=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(it):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts)), float(min(ts))


def micro():
    L = _lib.lib()
    sink = torch.zeros(4, device="cuda")
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    f2 = lambda mix: (lambda c, i, s, t: L.pigs_microbench_ffma2(c, i, mix, s, t))
    for name, f in (("ffma", L.pigs_microbench_ffma), ("ex2", L.pigs_microbench_ex2), ("ffma2", f2(0)), ("ffma2+mufu", f2(1))):
        for ctas_per_sm in (4, 8):
            ops = [0]

            def run():
                ops[0] = f(sms * ctas_per_sm, 20000, ctypes.c_void_p(sink.data_ptr()), st)
            med, best = ev_time(run)
            print(f"micro {name} ctas/sm={ctas_per_sm}: {ops[0] / best / 1e9:.1f} G lane-ops/s (best), "
                  f"{ops[0] / best / 1e6 / sms:.1f} lane-ops/ns/SM*1e-3", flush=True)


def perf(P, N, tag):
    rec = synthetic(P, N, seed=0)
    d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
    s = GaussianSampler(False, fused=False)
    med, best = ev_time(lambda: s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"]))
    pairs = s.pair_count()
    print(f"{tag}: P={P} N={N} R={s.num_rendered} pairs={pairs:.3e} preprocess {med:.3f} ms", flush=True)
    for mask in (1, 7):
        med, best = ev_time(lambda: s.sample_fused(mask))
        print(f"  fwd mask={mask}: {med:.3f} ms  {pairs / med / 1e6:.2f} Gpairs/s", flush=True)
    gs = [torch.randn(N, 1, device="cuda"), torch.randn(N, 2, 1, device="cuda"), torch.randn(N, 2, 2, 1, device="cuda"), None]
    from pigs_b200 import _C
    nr, bb, sbb, rg, srg = s._sample_state
    for mask in (1, 7):
        g = [gs[k] if mask & (1 << k) else None for k in range(4)]
        med, best = ev_time(lambda: _C.sample_backward(mask, d["means"], d["values"], d["conics"], d["samples"], nr, g, bb,
                                                       sbb, rg, srg, False))
        print(f"  bwd mask={mask}: {med:.3f} ms  {pairs / med / 1e6:.2f} Gpairs/s", flush=True)
    return s, d, pairs


_heap_set = False


def ref_perf(P, N):
    from tests import refcuda
    if not refcuda.available():
        print("reference CUDA not available")
        return
    # the reference allocates with device-side new[] in every thread (forward.cu:63-64, sampler_impl.cu:166):
    # raise the device malloc heap or it returns NULL beyond ~1e5 resident threads
    global _heap_set
    if not _heap_set:  # only legal before the first kernel that uses the device heap
        rt = ctypes.CDLL("libcudart.so.12")
        print("cudaDeviceSetLimit(MallocHeapSize, 8 GiB) ->", rt.cudaDeviceSetLimit(2, ctypes.c_size_t(8 << 30)))
        _heap_set = True
    rec = synthetic(P, N, seed=0)
    d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
    r = refcuda.RefSampler(False)
    med, _ = ev_time(lambda: r.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"]), 1, 3)
    print(f"REF: P={P} N={N} R={r.num_rendered} preprocess {med:.3f} ms", flush=True)
    for fn in range(3):
        med, _ = ev_time(lambda: r.forward(fn), 1, 2)
        print(f"  ref fwd fn={fn}: {med:.3f} ms", flush=True)
        g = torch.randn_like(r.forward(fn))
        med, _ = ev_time(lambda: r.backward(fn, g), 1, 2)
        print(f"  ref bwd fn={fn}: {med:.3f} ms", flush=True)


if __name__ == "__main__":
    print(torch.cuda.get_device_name(0))
    if "--ref" in sys.argv:
        # the unmodified reference CUDA extension (oracle/_ref/_C.so) on the same GPU, next to ours at the same size
        sizes = [(100_000, 1_000_000)] if "--ref-c2" in sys.argv else [(20_000, 200_000)]
        for P, N in sizes:
            ref_perf(P, N)
            perf(P, N, "OURS at the same size")
        sys.exit(0)
    if "--no-micro" not in sys.argv:
        micro()
    perf(100_000, 1_000_000, "C2")
    if "--big" in sys.argv:
        perf(1_000_000, 4_000_000, "C3")
    if "--shard8" in sys.argv:
        perf(1_000_000, 500_000, "C3 / 8 (one rank's shard at N = 8)")
