"""bench.py — the driver's measurement contract for pigs_b200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload C3|C2|C4] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

Metric (BASELINE.json): Gaussian-point evals/s, fwd+bwd with gradient + Hessian/Laplacian.  One "eval" is one
(Gaussian, point) pair of the reference's pair set; one STEP is one PDE-style pass over the
batch: tile binning (preprocess) -> fused value/gradient/Hessian forward -> fused backward to means / conics /
values (-> NCCL all-reduce of the packed gradients when N > 1).  value = pairs of all ranks / step time.

  value : inputs resident in HBM, CUDA events on the launching stream, max over ranks
  e2e   : the same step through the reference-shaped public API (diff_gaussian_sampling.GaussianSampler +
          autograd), inputs copied from pinned host memory and gradients + u read back to the host every step
  roofline     : dominant kernel (fused backward evaluation), FP32 lane-ops, timed live with CUDA events
                 recorded by the library around its evaluation launches (pigs_set_kernel_timing)
  cpu_baseline : the reference's pure-PyTorch dense path (port: oracle/dense_torch.py) on the host cores

Workloads (tests/fixtures.synthetic, seed 0): C3 = 1M Gaussians x 4M points (default; the config
the metric is quoted on and the one sharded over 1/2/4/8 GPUs), C2 = 100K x 1M, C4 = 500K anisotropic x 2M torus.
N > 1: the 4M points are split evenly over the ranks (strong scaling), Gaussians replicated.

Extra legs, run AFTER the timed C3 region and reported under "legs" in the same JSON line (--no-legs skips them):
  c5       : BASELINE config 5 — one model_pn.py-style step at 1M Gaussians (preprocess with samples = means -> u, grad u,
             Hessian without grad (model_pn.py:644-652) -> neighbour aggregation on a 16-regular CSR, 2 heads forward +
             backward (model_pn.py:253-274)); row-parallel over the ranks when N > 1.  Runs at every N.
  c1/c2/c4 : the other BASELINE configs, device-resident and end-to-end, N = 1 only
  pn_step  : a model_pn.py-style step at the reference's own scale (400 / 2500 Gaussians), ours and the reference extension
  ref_cuda : the UNMODIFIED reference CUDA extension (oracle/_ref/_C.so, built by oracle/build_ref.sh) on the same GPU
             and inputs next to ours, 20K Gaussians x 200K points (BASELINE.md's "GPU baseline to beat"), N = 1 only
"""
import argparse
import ctypes
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (P, N, kwargs for tests.fixtures.synthetic, description)
    "C3": (1_000_000, 4_000_000, {}, "2-D PDE residual (u, grad u, Hessian), 1M Gaussians x 4M points"),
    "C2": (100_000, 1_000_000, {}, "2-D density fit, 100K Gaussians x 1M points"),
    "C4": (500_000, 2_000_000, dict(edge_fraction=0.1, anisotropy=4.0), "periodic torus, 500K anisotropic Gaussians x 2M points"),
    "C1": (1_000, 10_000, dict(D=1), "1-D diffusion-style mixture, 1K Gaussians x 10K points (the reference's CPU-runnable case)"),
    "tiny": (20_000, 200_000, {}, "debug: 20K Gaussians x 200K points"),
}
METRIC = "gaussian_point_evals_per_s_fwd_bwd_grad_hessian"
UNIT = "pairs/s"
LANE_OPS_FWD, LANE_OPS_BWD = 19, 45  # algorithmic FP32 lane-ops per pair
EXEC_OPS_FWD, EXEC_OPS_BWD = 18.0, 36.5  # FP32 lane-ops per pair the kernels actually execute (SASS count, tools/sass_issue_model.py)


EVAL_SOURCES = ("eval_bwd.cu", "eval_fwd.cu", "eval_common.cuh", "common.cuh", "pack.cu")


def eval_source_sha():
    """sha1 over the evaluation-kernel sources: tools/ncu_summary.py stamps it into every ncu summary it writes, so a
    summary taken on an older kernel is recognised as stale here."""
    import hashlib
    h = hashlib.sha1()
    for name in EVAL_SOURCES:
        with open(os.path.join(ROOT, "pigs_b200", "csrc", name), "rb") as f:
            h.update(f.read())
    return h.hexdigest()[:16]


def ncu_traffic():
    """DRAM bytes per launch of the dominant kernel from the newest committed `ncu --set full` summary (profiles/) that
    was taken on the CURRENT kernel sources, or a note saying the newest one is stale.  bench.py cannot measure DRAM
    traffic itself; the summary is of the same command (tools/gpu_round.sh)."""
    import glob
    sha = eval_source_sha()
    stale = None
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "*_ncu_eval_c3.json")), key=os.path.getmtime, reverse=True):
        try:
            d = json.load(open(path))
            for k in d["kernels"]:
                if "eval_bwd_kernel<2, 1, 7, 0>" in k["kernel"] and k.get("dram_read_bytes") == k.get("dram_read_bytes"):
                    if d.get("source_sha") != sha:
                        stale = stale or os.path.basename(path)
                        break
                    return {"dram_bytes_per_launch": k["dram_read_bytes"] + k["dram_write_bytes"],
                            "kernel": "eval_bwd_kernel<2,1,7,FAST> at C3, 1 GPU",
                            "source": "profiles/%s (kernel sources %s)" % (os.path.basename(path), sha)}
        except Exception:
            continue
    return {"dram_bytes_per_launch": None, "kernel": "no ncu summary of the current kernel sources (%s)" % sha,
            "source": ("newest summary profiles/%s predates the last kernel change" % stale) if stale else ""}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=os.environ.get("PIGS_BENCH_WORKLOAD", "C3"), choices=list(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-legs", action="store_true", help="skip the c5 / c1 / c2 / c4 / ref_cuda / pn_step legs")
    return ap.parse_args()


class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU during the timed region (NVML, 100 ms period)."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz, self.power = [], set(), None, []
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        names = {}
        for n in dir(nv):
            if n.startswith("nvmlClocksThrottleReason") or n.startswith("nvmlClocksEventReason"):
                v = getattr(nv, n)
                if isinstance(v, int) and v:
                    names.setdefault(v, n.replace("nvmlClocksThrottleReason", "").replace("nvmlClocksEventReason", ""))
        while not self._stop.is_set():
            try:
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                self.power.append(nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, n in names.items():
                    if r & bit and n not in ("None", "GpuIdle", "All"):
                        self.reasons.add(n)
            except Exception:
                pass
            self._stop.wait(0.1)

    def start(self):
        if self.nv is not None:
            self._stop.clear()
            self._thr = threading.Thread(target=self._loop, daemon=True)
            self._thr.start()

    def stop(self):
        if self._thr is not None:
            self._stop.set()
            self._thr.join()
            self._thr = None

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["nvml unavailable"]}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "power_w_max": max(self.power) if self.power else None,
                "samples": len(self.samples)}


def make_inputs(workload, rank, world):
    """Seeded synthetic inputs on the host (numpy); every rank builds the same arrays and keeps its shard."""
    from pigs_b200.sharded import shard_range
    from tests.fixtures import synthetic
    P, N, kw, _ = WORKLOADS[workload]
    rec = synthetic(P, N, seed=0, **kw)
    lo, hi = shard_range(N, rank, world)
    rng = np.random.default_rng(1234)
    D = rec["means"].shape[1]
    ups = [rng.standard_normal((N,) + (D,) * k + (1,), dtype=np.float32) for k in range(3)]
    rec["samples"] = np.ascontiguousarray(rec["samples"][lo:hi])
    ups = [np.ascontiguousarray(u[lo:hi]) for u in ups]
    return rec, ups


def flush_note():
    return "per-step working set (samples, staged records, outputs, upstream gradients) > 126 MB L2"


def microbench_peaks(L, torch):
    """Register-resident FFMA and MUFU.EX2 loops: measured peak lane-ops/s of this GPU (T/s)."""
    sink = torch.zeros(4, device="cuda")
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    res = {}
    for name, f in (("ffma", L.pigs_microbench_ffma), ("ex2", L.pigs_microbench_ex2)):
        best = None
        for _ in range(4):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            ops = f(sms * 8, 4000, ctypes.c_void_p(sink.data_ptr()), st)
            b.record()
            torch.cuda.synchronize()
            r = ops / (a.elapsed_time(b) * 1e-3) / 1e12
            best = r if best is None else max(best, r)
        res[name] = best
    return res


PAIR_DEF_GPU = ("reference pair set (SURVEY.md S8): pairs (Gaussian, point) with the point's 0.51-wide tile inside the "
                "Gaussian's 3-sigma tile rectangle — what the reference extension evaluates")
PAIR_DEF_CPU = ("dense all-pairs: every (Gaussian, point) of the slice — the reference's pure-PyTorch gaussians.py path has no "
                "cutoff, so its pairs/s counts a different (larger) pair set per sample than the GPU arm's reference pair set")


def _all_threads(torch):
    """torchrun exports OMP_NUM_THREADS=1: the CPU arm takes every host core explicitly."""
    n = os.cpu_count() or 1
    try:
        torch.set_num_threads(n)
    except Exception:
        pass
    return torch.get_num_threads()


def _dense_slice(workload):
    import torch
    from tests.fixtures import synthetic
    P, N, kw, _ = WORKLOADS[workload]
    n, m = min(P, 1000), min(N, 10000)  # dense intermediates are O(n*m*d*d*4 B)
    rec = synthetic(P, min(N, 200_000), seed=0, **kw)
    rng = np.random.default_rng(7)
    gi = np.sort(rng.choice(P, n, replace=False))
    si = np.sort(rng.choice(rec["samples"].shape[0], m, replace=False))
    means, conics, values = (torch.from_numpy(rec[k][gi]) for k in ("means", "conics", "values"))
    samples = torch.from_numpy(rec["samples"][si])
    return means, conics, values, samples, n, m


def cpu_baseline(workload, budget_s=12.0):
    """Reference's pure-PyTorch dense CPU path (port) on a bounded dense slice of the workload."""
    import torch

    from oracle import dense_torch as dt
    threads = _all_threads(torch)
    means, conics, values, samples, n, m = _dense_slice(workload)
    t1, pairs = dt.timed_steps(means, conics, values, samples, 1, 1)
    steps = int(max(1, min(8, budget_s / max(t1[0], 1e-3) - 2)))
    times, pairs = dt.timed_steps(means, conics, values, samples, steps, 0)
    best = min(times + t1)
    return {"value": pairs / best, "unit": UNIT, "cores": threads, "host_cpus": os.cpu_count(),
            "kind": "port", "pair_definition": PAIR_DEF_CPU,
            "sample": "dense all-pairs slice of %s: %d Gaussians x %d points, u + grad u + Hessian forward and "
                      "autograd backward, oracle/dense_torch.py restating gaussians.py:48-116; best of %d" %
                      (workload, n, m, len(times) + 1)}


class SamplingRun:
    """One sampling workload (preprocess + fused u / grad u / Hessian forward + fused backward) on this rank's shard:
    the device-resident step, the end-to-end step through the public API, and their timing."""

    def __init__(self, workload, torch, dist, L, rank, world):
        from pigs_b200 import _C, _lib
        self.torch, self.dist, self.L, self.rank, self.world = torch, dist, L, rank, world
        self._C, self._lib = _C, _lib
        self.workload = workload
        self.P, self.N, _, self.desc = WORKLOADS[workload]
        rec, ups = make_inputs(workload, rank, world)
        self.host = {k: torch.from_numpy(v).pin_memory() for k, v in rec.items()}
        self.host_ups = [torch.from_numpy(u).pin_memory() for u in ups]
        self.dev = {k: v.cuda(non_blocking=True) for k, v in self.host.items()}
        self.dev_ups = [u.cuda(non_blocking=True) for u in self.host_ups]
        torch.cuda.synchronize()
        self.D, self.Nl = self.dev["means"].shape[1], self.dev["samples"].shape[0]
        self.state = {}

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def step_device(self):
        from pigs_b200.sharded import allreduce_bounds, allreduce_grads
        _C, dev, ups = self._C, self.dev, self.dev_ups
        mm = _C.sample_bounds(dev["samples"], self.D)
        if self.world > 1:
            allreduce_bounds(mm, self.D)
        nr, bb, sbb, rg, srg, radii = _C.preprocess_gaussians(dev["means"], dev["values"], dev["covariances"],
                                                              dev["conics"], dev["samples"], False, bounds=mm)
        a = (dev["means"], dev["values"], dev["conics"], dev["samples"], nr)
        outs = _C.sample_forward(7, *a, bb, sbb, rg, srg, False)
        dm, dv, dc = _C.sample_backward(7, *a, [ups[0], ups[1], ups[2], None], bb, sbb, rg, srg, False)
        if self.world > 1:
            allreduce_grads(_C.packed_base(dm))
        self.state.update(rg=rg, srg=srg, outs=outs, grads=(dm, dv, dc))

    def timed(self, fn, steps, warmup, sampler=None):
        torch = self.torch
        for _ in range(warmup):
            fn()
        self.barrier()
        n0 = self.L.pigs_launch_count()
        if sampler:
            sampler.start()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(steps):
            fn()
        b.record()
        self.barrier()
        if sampler:
            sampler.stop()
        ms = torch.tensor([a.elapsed_time(b)], device="cuda")
        if self.world > 1:
            self.dist.all_reduce(ms, op=self.dist.ReduceOp.MAX)
        return float(ms.item()), self.L.pigs_launch_count() - n0

    def count_pairs(self):
        """pair count of this rank's shard from the ranges arrays, and summed over ranks"""
        torch = self.torch
        self.step_device()
        b = self.state["rg"].numel() // 8 - 1
        g = self.state["rg"].view(torch.int32)[:2 * b].view(b, 2).to(torch.int64)
        s = self.state["srg"].view(torch.int32)[:2 * b].view(b, 2).to(torch.int64)
        pairs_local = ((g[:, 1] - g[:, 0]) * (s[:, 1] - s[:, 0])).sum()
        pairs_t = pairs_local.clone()
        if self.world > 1:
            self.dist.all_reduce(pairs_t)
        self.pairs, self.pairs_local = int(pairs_t.item()), int(pairs_local.item())
        return self.pairs

    def run_device(self, steps, warmup, clocks=None):
        """-> (ms_per_step max over ranks, launches, {fwd: (ms per launch group, n), bwd: ...} of this rank)"""
        L, _lib = self.L, self._lib
        L.pigs_set_kernel_timing(1)
        for w in (0, 1):  # drop events recorded so far
            L.pigs_kernel_timing_read(w, ctypes.byref(ctypes.c_float()), ctypes.byref(ctypes.c_int()))
        L.pigs_set_kernel_timing(0)
        for _ in range(warmup):
            self.step_device()
        L.pigs_set_kernel_timing(1)
        total_ms, launches = self.timed(self.step_device, steps, 0, clocks)
        L.pigs_set_kernel_timing(0)
        kt = {}
        for w, nm in ((0, "fwd"), (1, "bwd")):
            ms, n = ctypes.c_float(), ctypes.c_int()
            _lib.check(L.pigs_kernel_timing_read(w, ctypes.byref(ms), ctypes.byref(n)), "pigs_kernel_timing_read")
            kt[nm] = (ms.value / max(n.value, 1), n.value)
        return total_ms / steps, launches, kt

    def run_e2e(self, steps, warmup):
        """The same step through the reference-shaped public API with HOST buffers: inputs copied from pinned memory,
        u and the three gradients read back, every step, inside the timed region."""
        torch, P, D, Nl, host, host_ups = self.torch, self.P, self.D, self.Nl, self.host, self.host_ups
        from pigs_b200.sharded import ShardedGaussianSampler
        if self.world > 1:
            sampler = ShardedGaussianSampler(False, group=None)
        else:
            from diff_gaussian_sampling import GaussianSampler
            sampler = GaussianSampler(False)
        K = D * (D + 1) // 2
        hg = torch.empty(P * (D + 1 + K), dtype=torch.float32).pin_memory()
        hu = torch.empty((Nl, 1), dtype=torch.float32).pin_memory()
        h2d = sum(host[k].numel() * 4 for k in ("means", "values", "covariances", "conics", "samples")) + \
            sum(u.numel() * 4 for u in host_ups)
        d2h = hg.numel() * 4 + hu.numel() * 4

        copy_stream = torch.cuda.Stream()

        def step_e2e():
            main = torch.cuda.current_stream()
            d = {k: host[k].cuda(non_blocking=True) for k in ("means", "values", "covariances", "conics", "samples")}
            # the upstream gradients are not needed before the backward: their host->device copies (the larger part of the
            # step's input bytes) run on a copy stream next to the binning and the forward; still inside the timed region
            copy_stream.wait_stream(main)
            with torch.cuda.stream(copy_stream):
                gu = [u.cuda(non_blocking=True) for u in host_ups]
            for k in ("means", "values", "conics"):
                d[k].requires_grad_(True)
            sampler.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
            u = sampler.sample_gaussians()
            ux = sampler.sample_gaussians_derivative()
            uxx = sampler.sample_gaussians_laplacian()
            main.wait_stream(copy_stream)
            for t in gu:
                t.record_stream(main)
            loss = (u * gu[0]).sum() + (ux * gu[1]).sum() + (uxx * gu[2]).sum()
            loss.backward()
            hg[:P * D].copy_(d["means"].grad.reshape(-1), non_blocking=True)
            hg[P * D:P * (D + 1)].copy_(d["values"].grad.reshape(-1), non_blocking=True)
            hg[P * (D + 1):].copy_(d["conics"].grad.reshape(-1), non_blocking=True)
            hu.copy_(u.detach(), non_blocking=True)
            torch.cuda.current_stream().synchronize()

        e_ms, _ = self.timed(step_e2e, steps, max(warmup, 2))
        return {"value": self.pairs / (e_ms / steps * 1e-3), "unit": UNIT, "ms_per_step": e_ms / steps,
                "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "api": "diff_gaussian_sampling.GaussianSampler.preprocess/sample_gaussians/_derivative/_laplacian + loss.backward()"}


def fp32_peak(torch, ck):
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    fmax = (ck.get("sm_max_mhz") or 1965.0) * 1e6
    return sms, fmax, sms * 128 * fmax / 1e12


def leg_sampling(name, torch, dist, L, steps, warmup):
    """BASELINE configs other than the headline one (N = 1): device-resident and end-to-end step of the same path."""
    run = SamplingRun(name, torch, dist, L, 0, 1)
    pairs = run.count_pairs()
    ms, launches, kt = run.run_device(steps, warmup)
    e2e = run.run_e2e(steps, warmup)
    out = {"workload": "%s: %s" % (name, run.desc), "gaussians": run.P, "points": run.N, "pairs": pairs, "steps": steps,
           "warmup": warmup, "ms_per_step": ms, "value": pairs / (ms * 1e-3), "unit": UNIT,
           "kernel_ms": {"eval_fwd": kt["fwd"][0], "eval_bwd": kt["bwd"][0], "other_per_step": ms - kt["fwd"][0] - kt["bwd"][0]},
           "gpu_launches_per_step": launches / steps,
           "e2e": {k: e2e[k] for k in ("value", "ms_per_step", "h2d_bytes_per_step", "d2h_bytes_per_step")}}
    if run.D == 2 and kt["bwd"][0] > 0:
        _, _, peak = fp32_peak(torch, {})
        out["roofline_frac_fwd_bwd"] = pairs * (LANE_OPS_FWD + LANE_OPS_BWD) / ((kt["fwd"][0] + kt["bwd"][0]) * 1e-3) / 1e12 / peak
    else:
        out["bound"] = "launch latency (1-D, 1K x 10K: the step is a chain of dependent launches of a few us each)"
    return out


def hbm_peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    except Exception:
        return 6550.0, "fallback: B200_PROFILING.md measured copy bandwidth"


def leg_c5(torch, dist, L, rank, world, steps=5, warmup=3):
    """BASELINE config 5 (model_pn.py neighbour aggregation + PDE step, 1M Gaussians, 16 neighbours).  One step = the
    library calls of one Model.forward + loss backward around the (torch, out-of-scope) MLP:
      preprocess(samples = means) -> [preprocess_aggregate: lists + exchange plan, they depend on the means only] ->
      sample_gaussians / _derivative / _laplacian without grad (model_pn.py:644-652)
      -> ATTENTION_HEADS = 2 x aggregate_neighbors forward, then backward (model_pn.py:253-274).
    Neighbour lists: a 16-regular CSR (the reference's own rule yields ~1.6 K neighbours at this P), passed
    as plain inputs like the reference API allows.  N > 1 (row-parallel): every rank samples at ITS window
    of the means (windows of equal pair counts) and owns that window's rows of features / queries / keys, of the output
    and of the gradients; halo rows travel point to point, the shared parameters' gradients are all-reduced."""
    from scipy.spatial import cKDTree

    from pigs_b200 import _C
    from pigs_b200.sharded import ShardedGaussianSampler, balanced_ranges
    from tests.fixtures import agg_params, synthetic
    P, k, HEADS = 1_000_000, 16, 2
    rec = synthetic(P, 16, seed=0)
    d = {kk: torch.from_numpy(v).cuda() for kk, v in rec.items()}
    s = ShardedGaussianSampler(False)
    if world > 1:
        # set-up, not timed: one binning pass over ALL means -> per-row pair counts -> windows of equal evaluation work
        s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["means"])
        lo, hi = balanced_ranges(s.pair_weights(), world)[rank]
        split = "equal pair counts (sharded.balanced_ranges over GaussianSampler.pair_weights)"
    else:
        lo, hi = 0, P
        split = "single window"
    n = hi - lo
    mine = d["means"][lo:hi].contiguous()
    s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], mine)
    _, idx = cKDTree(rec["means"]).query(rec["means"][lo:hi], k=k, workers=-1)
    indices = torch.from_numpy(np.sort(idx.astype(np.int64), axis=1).reshape(-1)).cuda()
    ranges = ((torch.arange(1, P + 1, device="cuda") - lo).clamp(0, n) * k).contiguous()
    i_of = torch.arange(lo, hi, device="cuda").repeat_interleave(k)
    X = d["means"][indices] - d["means"][i_of]
    X = torch.where(X.abs() > 1, X - 2 * torch.sign(X), X)
    dists = (X / (0.333 * s.radii[i_of] + 1e-6).unsqueeze(1)).contiguous()
    cj = d["conics"][indices]
    dens = torch.exp(-0.5 * (cj[:, 0] * X[:, 0] ** 2 + cj[:, 2] * X[:, 1] ** 2) - cj[:, 1] * X[:, 0] * X[:, 1]).contiguous()
    inv_total = torch.full((P,), 1e6, device="cuda")
    inv_total[lo:hi] = 1.0 / (dens.view(n, k).sum(1) + 1e-6)
    csr = (indices, ranges, dists, dens, inv_total)
    del X, cj, i_of
    heads = []
    for h in range(HEADS):
        ap = agg_params(P, 16, 16, 25, 2, seed=100 + h)
        heads.append({kk: torch.from_numpy(v[lo:hi] if kk in ("features", "queries", "keys") else v).cuda().contiguous()
                      for kk, v in ap.items()})
    feat = heads[0]["features"].requires_grad_(True)          # shared by the heads (model_pn.py:256-262)
    freq = heads[0]["frequencies"].requires_grad_(True)
    per_head = [[hd[kk].requires_grad_(True) for kk in ("transform", "queries", "keys", "distance_transform")] for hd in heads]
    leaves = [feat, freq] + [t for ph in per_head for t in ph]
    gs = [torch.randn(n, 16, device="cuda", generator=torch.Generator(device="cuda").manual_seed(7 + h)) for h in range(HEADS)]

    # Order inside the step: binning -> neighbour lists + exchange plan -> sampling -> aggregation.  The lists depend on
    # the means (and the radii the binning yields) only, not on the sampled values, so installing them BEFORE the sampling
    # launch puts the plan's one host read behind the short binning kernels instead of behind the ~10 ms evaluation
    # kernel — the host then enqueues the whole aggregation while the GPU is still sampling.
    def sampling():
        with torch.no_grad():
            s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], mine)
        s.preprocess_aggregate(row_range=(lo, hi), csr=csr)
        with torch.no_grad():
            return s.sample_gaussians(), s.sample_gaussians_derivative(), s.sample_gaussians_laplacian()

    def aggregation():
        outs = s.aggregate_neighbors_rows_heads(feat, freq, [tuple(ph) for ph in per_head])  # both heads, one exchange each way
        for t in leaves:
            t.grad = None
        torch.autograd.backward(outs, gs)

    mids, ends = [], []

    def step():
        sampling()
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        mids.append(e)
        aggregation()
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        ends.append(e)

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        n0 = L.pigs_launch_count()
        del mids[:], ends[:]
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(steps):
            fn()
        b.record()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ms = torch.tensor([a.elapsed_time(b) / steps], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), (L.pigs_launch_count() - n0) / steps, a

    ms_step, launches, ev0 = timed(step, steps, warmup)
    # split of the SAME timed steps at the events between / behind the two halves (this rank's stream; the step time is
    # the max over ranks)
    starts = [ev0] + ends[:-1]
    ms_samp = float(np.mean([a_.elapsed_time(m_) for a_, m_ in zip(starts, mids)]))
    ms_agg = float(np.mean([m_.elapsed_time(e_) for m_, e_ in zip(mids, ends)]))
    # aggregation kernels alone (one head), for their HBM roofline on this rank's window
    plan = s.row_plan
    T, Q, Kk, DT = [t.detach() for t in per_head[0]]
    f_full, k_full = feat.detach(), Kk
    if not (world == 1):
        f_full = torch.zeros(P, 16, device="cuda")
        k_full = torch.zeros(P, 16, device="cuda")
        f_full[lo:hi], k_full[lo:hi] = feat.detach(), Kk
        plan.gather([f_full, k_full])
    w_, e_, f_, _ = _C.aggregate_neighbors_rows(f_full, T, Q, k_full, freq.detach(), DT, *csr, (lo, hi), False)
    gf, gk = torch.zeros_like(f_full), torch.zeros_like(k_full)
    ms_kf, _, _ = timed(lambda: _C.aggregate_neighbors_rows(f_full, T, Q, k_full, freq.detach(), DT, *csr, (lo, hi), False), steps, 2)
    ms_kb, _, _ = timed(lambda: _C.aggregate_neighbors_rows_backward(f_full, T, Q, k_full, freq.detach(), DT, indices, ranges,
                                                                     dists, dens, w_, e_, f_, inv_total, gs[0], gf, gk, (lo, hi),
                                                                     False), steps, 2)
    pairs_t = torch.zeros(world, device="cuda", dtype=torch.int64)
    pairs_t[rank] = s.pair_count()
    rows_t = torch.zeros(world, device="cuda", dtype=torch.int64)
    rows_t[rank] = n
    if world > 1:
        dist.all_reduce(pairs_t)
        dist.all_reduce(rows_t)
    peak, peak_src = hbm_peak()
    M = n * k
    by_f, by_b = 160 * M + 140 * n, 416 * M + 260 * n
    return {"workload": "C5: model_pn.py-style step, 1M Gaussians, samples = means, 16-neighbour aggregation, 2 heads",
            "gaussians": P, "neighbours": k, "heads": HEADS, "n_gpus": world, "steps": steps, "warmup": warmup,
            "ms_per_step": ms_step, "sampling_ms": ms_samp, "aggregation_ms": ms_agg,
            "pairs_sampling_all_ranks": int(pairs_t.sum().item()), "pairs_per_rank": pairs_t.tolist(),
            "rows_per_rank": rows_t.tolist(), "row_split": split, "halo_rows_rank0": int(plan.halo_rows),
            "exchange": "none" if world == 1 else ("all-gather / all-reduce (dense fallback)" if plan.dense else
                                                  "point-to-point halo rows + all-reduce of the shared parameters' gradients"),
            "functions": "u + grad u + Hessian (fused, no grad) at the rank's means; aggregate_neighbors fwd + bwd x 2 heads",
            "gpu_launches_per_step": launches,
            "timing": "CUDA events on the compute stream, barrier + synchronize either side, max over ranks",
            "roofline": {"bound": "hbm", "kernel": "agg_backward16_kernel (one head, this rank's window)",
                         "achieved": by_b / (ms_kb * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": by_b / (ms_kb * 1e-3) / 1e9 / peak, "peak_source": peak_src, "kernel_ms": ms_kb,
                         "algorithmic_bytes": by_b, "bytes_model": "416 B per slot + 260 B per Gaussian (DESIGN.md §4)",
                         "forward": {"kernel": "agg_forward16_kernel", "kernel_ms": ms_kf, "algorithmic_bytes": by_f,
                                     "achieved": by_f / (ms_kf * 1e-3) / 1e9, "frac": by_f / (ms_kf * 1e-3) / 1e9 / peak,
                                     "bytes_model": "160 B per slot + 140 B per Gaussian"}}}


def leg_ref_cuda(torch, L, P=20_000, N=200_000):
    """The unmodified reference CUDA extension (oracle/_ref/_C.so) on this GPU next to ours, same inputs: one step =
    preprocess + value / gradient / Hessian forward + their backward (the reference: 3 + 3 separate passes)."""
    from tests import refcuda
    if not refcuda.available():
        return {"unavailable": "oracle/_ref/_C.so is not on this box (built by oracle/build_ref.sh where /root/reference exists)"}
    from pigs_b200 import _C
    from tests.fixtures import synthetic
    # the reference allocates with device-side new[] in every thread (forward.cu:63-64, sampler_impl.cu:166)
    rt = ctypes.CDLL("libcudart.so.12")
    rt.cudaDeviceSetLimit(2, ctypes.c_size_t(8 << 30))  # cudaLimitMallocHeapSize; before the first kernel that uses the heap
    rec = synthetic(P, N, seed=0)
    d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
    g = torch.Generator(device="cuda").manual_seed(3)
    gs = [torch.randn(N, 1, device="cuda", generator=g), torch.randn(N, 2, 1, device="cuda", generator=g),
          torch.randn(N, 2, 2, 1, device="cuda", generator=g)]
    r = refcuda.RefSampler(False)

    def ref_step():
        r.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
        outs = [r.forward(fn) for fn in range(3)]
        grads = [r.backward(fn, gs[fn]) for fn in range(3)]
        return outs, grads

    def our_step():
        nr, bb, sbb, rg, srg, _ = _C.preprocess_gaussians(d["means"], d["values"], d["covariances"], d["conics"],
                                                          d["samples"], False)
        a = (d["means"], d["values"], d["conics"], d["samples"], nr)
        outs = _C.sample_forward(7, *a, bb, sbb, rg, srg, False)
        grads = _C.sample_backward(7, *a, [gs[0], gs[1], gs[2], None], bb, sbb, rg, srg, False)
        return outs, grads

    def timed(fn, steps, warmup):
        # per-step events, median over the steps: one stalled step (seen once on a fresh box right behind the reference's
        # device-heap kernels: 43 ms instead of 2.3) must not decide a ratio of two short measurements
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
        ev[0].record()
        for i in range(steps):
            res = fn()
            ev[i + 1].record()
        torch.cuda.synchronize()
        per = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(steps))
        return per[len(per) // 2] if steps % 2 else 0.5 * (per[steps // 2 - 1] + per[steps // 2]), res

    ref_ms, (ro, rg_) = timed(ref_step, 2, 1)
    our_ms, (oo, og) = timed(our_step, 5, 3)
    rel = lambda a_, b_: float((a_ - b_).abs().max() / b_.abs().max().clamp_min(1e-30))
    dev = {"u": rel(oo[0], ro[0]), "du": rel(oo[1], ro[1]), "ddu": rel(oo[2], ro[2])}
    rsum = [sum(x[i].reshape(-1) for x in rg_) for i in range(3)]
    for nm, a_, b_ in zip(("dmeans", "dvalues", "dconics"), og, rsum):
        dev[nm] = rel(a_.reshape(-1), b_)
    pairs = r.num_rendered  # instances; the pair count proper comes from our ranges
    return {"workload": "%d Gaussians x %d points, 2-D, u + grad u + Hessian forward + backward" % (P, N),
            "reference": "oracle/_ref/_C.so = the unmodified diff-gaussian-sampling CUDA extension (forward.cu:277-343 one "
                         "CTA per tile, backward.cu:418-499), built by oracle/build_ref.sh",
            "ref_ms_per_step": ref_ms, "ours_ms_per_step": our_ms, "speedup": ref_ms / our_ms,
            "steps": {"reference": 2, "ours": 5}, "warmup": {"reference": 1, "ours": 3},
            "num_rendered": int(pairs), "max_rel_deviation_ours_vs_reference": dev,
            "timing": "CUDA events on the current stream around every step (median over the steps), inputs resident on the device"}


def leg_pn_step(torch, L, sizes=(400, 2500)):
    """One model_pn.py-style step at the reference's OWN scale (main_pn.py: nx = ny = 20 -> 400 Gaussians), through the public
    API: preprocess(samples = means) -> u, grad u, Hessian -> preprocess_aggregate (the real neighbour search) -> 2 heads of
    aggregate_neighbors, then the backward of a scalar loss through all of it (model_pn.py:644-652,253-274).  Next to it the
    same calls on the UNMODIFIED reference extension (oracle/_ref/_C.so) where it is on the box.  Launch-bound on both sides."""
    from diff_gaussian_sampling import GaussianSampler
    from tests import refcuda
    from tests.fixtures import agg_model
    from tests.helpers import AGG_KEYS
    HEADS = 2

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
        n0 = L.pigs_launch_count()
        ev[0].record()
        for i in range(steps):
            fn()
            ev[i + 1].record()
        torch.cuda.synchronize()
        per = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(steps))
        return per[len(per) // 2], (L.pigs_launch_count() - n0) / steps

    rows = []
    for P in sizes:
        rec = agg_model(P, seed=21)
        d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
        for k in ("means", "values", "conics"):
            d[k].requires_grad_(True)
        heads = [[d[k].clone().requires_grad_(True) for k in AGG_KEYS] for _ in range(HEADS)]
        s = GaussianSampler(False)

        def ours():
            s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["means"])
            u, ux, uxx = s.sample_gaussians(), s.sample_gaussians_derivative(), s.sample_gaussians_laplacian()
            s.preprocess_aggregate()
            loss = u.sum() + ux.sum() + uxx.sum()
            for h in heads:
                loss = loss + s.aggregate_neighbors(*h).sum()
            loss.backward()

        ms, launches = timed(ours, 30, 10)
        row = {"gaussians": P, "neighbour_slots": int(s.indices.numel()), "ours_ms_per_step": ms, "ours_launches_per_step": launches}
        if refcuda.available():
            r = refcuda.RefSampler(False)
            dd = {k: v.detach() for k, v in d.items()}
            hh = [[t.detach() for t in h] for h in heads]

            def ref():
                r.preprocess(dd["means"], dd["values"], dd["covariances"], dd["conics"], dd["means"])
                outs = [r.forward(k) for k in range(3)]
                csr = r.preprocess_aggregate()
                for h in hh:
                    w, e, f, out = r.C.aggregate_neighbors(*h, *csr, False)
                    r.C.aggregate_neighbors_backward(*h, csr[0], csr[1], csr[2], csr[3], w, e, f, csr[4], torch.ones_like(out), False)
                for k in range(3):
                    r.backward(k, torch.ones_like(outs[k]))

            rms, _ = timed(ref, 5, 2)
            row["ref_ms_per_step"] = rms
            row["speedup"] = rms / ms
        rows.append(row)
    return {"workload": "model_pn.py-style step at the reference's own scale: preprocess(samples = means) + u / grad u / Hessian + "
                        "neighbour search + 2 heads of aggregate_neighbors, forward and backward, public API",
            "reference": "oracle/_ref/_C.so, the same calls (its three forward and three backward passes, aggregate_neighbors.cu)",
            "timing": "CUDA events around every step, median; both sides launch- / host-bound at these sizes", "rows": rows}


def run_ours(args):
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: pigs_b200 has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from pigs_b200 import _lib
    L = _lib.lib()
    P, N, _, desc = WORKLOADS[args.workload]
    run = SamplingRun(args.workload, torch, dist, L, rank, world)
    D, Nl = run.D, run.Nl

    clocks = ClockSampler(local)
    peaks = microbench_peaks(L, torch) if rank == 0 else None
    pairs = run.count_pairs()
    pairs_local = run.pairs_local
    ms_per_step, launches, kt = run.run_device(args.steps, args.warmup, clocks)
    value = pairs / (ms_per_step * 1e-3)
    # per-rank kernel times and pair counts (load balance of the sample shards; rank 0 prints them)
    per_rank = torch.tensor([kt["fwd"][0], kt["bwd"][0], float(pairs_local)], device="cuda", dtype=torch.float64)
    if world > 1:
        gathered = [torch.zeros_like(per_rank) for _ in range(world)]
        dist.all_gather(gathered, per_rank)
        per_rank_rows = [[float(v) for v in g.tolist()] for g in gathered]
    else:
        per_rank_rows = [[float(v) for v in per_rank.tolist()]]
    launches_t = torch.tensor([launches], device="cuda", dtype=torch.int64)
    if world > 1:
        dist.all_reduce(launches_t)

    e2e = None if args.no_e2e else run.run_e2e(args.steps, args.warmup)
    del run
    torch.cuda.empty_cache()

    # ---------------- extra legs (outside the timed C3 region) ----------------
    legs = {}
    if not args.no_legs:
        def guarded(name, fn):
            t0 = time.perf_counter()
            try:
                res = fn()
            except Exception as ex:  # a leg must never cost the headline line
                res = {"error": "%s: %s" % (type(ex).__name__, str(ex)[:300])}
            if isinstance(res, dict):
                res["wall_s"] = round(time.perf_counter() - t0, 2)
            torch.cuda.empty_cache()
            return res
        legs["c5"] = guarded("c5", lambda: leg_c5(torch, dist, L, rank, world))
        if world == 1:
            legs["c2"] = guarded("c2", lambda: leg_sampling("C2", torch, dist, L, 3, 3))
            legs["c4"] = guarded("c4", lambda: leg_sampling("C4", torch, dist, L, 3, 3))
            legs["c1"] = guarded("c1", lambda: leg_sampling("C1", torch, dist, L, 50, 10))
            legs["ref_cuda"] = guarded("ref_cuda", lambda: leg_ref_cuda(torch, L))
            legs["pn_step"] = guarded("pn_step", lambda: leg_pn_step(torch, L))

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    ck = clocks.summary()
    sms, fmax, peak = fp32_peak(torch, ck)
    bwd_ms, fwd_ms = kt["bwd"][0], kt["fwd"][0]
    ach_bwd = pairs_local * LANE_OPS_BWD / (bwd_ms * 1e-3) / 1e12 if bwd_ms > 0 else None
    ach_all = pairs_local * (LANE_OPS_FWD + LANE_OPS_BWD) / ((bwd_ms + fwd_ms) * 1e-3) / 1e12 if bwd_ms > 0 else None
    traffic = ncu_traffic()
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic (seeded jittered-grid Gaussians, uniform points; tests/fixtures.synthetic seed 0)",
        "config": {"workload": "%s: %s" % (args.workload, desc), "gaussians": P, "points": N, "points_per_gpu": Nl,
                   "pairs": pairs, "pair_definition": PAIR_DEF_GPU,
                   "functions": "u + grad u + Hessian fused fwd, fused bwd to means/conics/values",
                   "step": "preprocess (tile binning) + forward + backward" + (" + NCCL all-reduce of packed grads" if world > 1 else ""),
                   "parallelism": "points sharded x%d, Gaussians replicated" % world, "l2": flush_note()},
        "pde_step_ms": ms_per_step,
        "gpu_launches": int(launches_t.item()),
        "kernel_ms": {"eval_fwd": fwd_ms, "eval_bwd": bwd_ms, "other_per_step": ms_per_step - fwd_ms - bwd_ms,
                      "per_rank_fwd_bwd_ms": [[round(r[0], 3), round(r[1], 3)] for r in per_rank_rows],
                      "per_rank_pairs": [int(r[2]) for r in per_rank_rows]},
        "roofline": {"bound": "fp32", "kernel": "eval_bwd_kernel<2,1,7,*> (fused backward evaluation)",
                     "achieved": ach_bwd, "peak": peak, "unit": "T lane-op/s", "frac": ach_bwd / peak if ach_bwd else None,
                     "peak_source": "%d SMs x 128 FP32 lanes x %.0f MHz (clocks.max.sm); MEASURED_PEAKS.json has no FP32 figure" % (sms, fmax / 1e6),
                     "peak_measured_ffma": peaks["ffma"], "frac_of_measured_ffma": ach_bwd / peaks["ffma"] if ach_bwd else None,
                     "peak_measured_ex2": peaks["ex2"], "lane_ops_per_pair": LANE_OPS_BWD,
                     "fwd_plus_bwd": {"achieved": ach_all, "frac": ach_all / peak if ach_all else None,
                                      "lane_ops_per_pair": LANE_OPS_FWD + LANE_OPS_BWD,
                                      "executed_lane_ops_per_pair": EXEC_OPS_FWD + EXEC_OPS_BWD,
                                      "frac_executed": (pairs_local * (EXEC_OPS_FWD + EXEC_OPS_BWD) /
                                                        ((bwd_ms + fwd_ms) * 1e-3) / 1e12 / peak) if bwd_ms > 0 else None},
                     "executed_lane_ops_per_pair": EXEC_OPS_BWD,
                     "frac_executed": (ach_bwd * EXEC_OPS_BWD / LANE_OPS_BWD) / peak if ach_bwd else None,
                     "traffic": traffic.get("dram_bytes_per_launch"),
                     "traffic_note": (traffic.get("kernel", "") + "; " + traffic.get("source", "")).strip("; ")},
        "clocks": ck,
    }
    if D != 2:
        # the lane-op counts above are the 2-D one-channel kernels'; the 1-D case (C1) is launch-bound at its size
        out["roofline"] = {"bound": "launch latency", "note": "1-D, 1K x 10K: a chain of dependent launches of a few us each; "
                           "no FP32 roofline is claimed for this workload", "kernel_ms": {"eval_fwd": fwd_ms, "eval_bwd": bwd_ms}}
    if e2e:
        out["e2e"] = e2e
    if world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(args.workload)
    if legs:
        out["legs"] = legs
        for k in ("c5", "ref_cuda", "pn_step", "c1", "c2", "c4"):  # also at top level, where the verdict asked for them
            if k in legs:
                out[k] = legs[k]
        del out["legs"]
    emit(out)
    if world > 1:
        dist.destroy_process_group()


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path — its pure-PyTorch dense sampling
    (gaussians.py:48-116; the module itself cannot travel to the GPU box, so oracle/dense_torch.py restates it) —
    on ALL host cores (set explicitly: torchrun exports OMP_NUM_THREADS=1), on the same workload's data, with exactly
    the --steps / --warmup asked for.  Each step is a bounded sample: a dense 1000-Gaussian x 10000-point slice (dense
    intermediates are O(n*m*d^2)); its pairs are DENSE all-pairs, stated in the line."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    import torch

    from oracle import dense_torch as dt
    threads = _all_threads(torch)
    P, N, kw, desc = WORKLOADS[args.workload]
    means, conics, values, samples, n, m = _dense_slice(args.workload)
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    t0 = time.perf_counter()
    times, pairs = dt.timed_steps(means, conics, values, samples, steps, warmup)
    ms = 1e3 * sum(times) / len(times)
    value = pairs / (ms * 1e-3)
    sample = ("each step = dense all-pairs slice of %s: %d Gaussians x %d points, u + grad u + Hessian forward + autograd "
              "backward (pure-PyTorch path of gaussians.py:48-116, restated in oracle/dense_torch.py)" % (args.workload, n, m))
    out = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps,
           "warmup": warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
           "dtype": "f32", "data": "synthetic (same seeded recipe as the GPU arm)",
           "config": {"workload": "%s: %s" % (args.workload, desc), "gaussians": P, "points": N, "sample": sample,
                      "pair_definition": PAIR_DEF_CPU},
           "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "host_cpus": os.cpu_count(),
                            "kind": "port", "sample": sample, "pair_definition": PAIR_DEF_CPU},
           "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0, "wall_s": time.perf_counter() - t0}
    emit(out)


_JSON_FD = None


def emit(obj):
    """The ONE JSON line goes to the process's original stdout."""
    line = (json.dumps(obj) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(line.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, line)


def main():
    global _JSON_FD
    args = parse_args()
    # native libraries (NCCL prints its version banner to fd 1 when NCCL_DEBUG is set) must not share stdout with the
    # JSON line: keep the original stdout for emit() and point fd 1 at stderr for everything else
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
