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


def ncu_traffic():
    """DRAM bytes per launch of the dominant kernel from the newest committed `ncu --set full` summary (profiles/), or
    None.  bench.py cannot measure DRAM traffic itself; the summary is of the same command (tools/gpu_round.sh)."""
    import glob
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "*_ncu_eval_c3.json")), reverse=True):
        try:
            d = json.load(open(path))
            for k in d["kernels"]:
                if "eval_bwd_kernel<2, 1, 7, 0>" in k["kernel"] and k.get("dram_read_bytes") == k.get("dram_read_bytes"):
                    return {"dram_bytes_per_launch": k["dram_read_bytes"] + k["dram_write_bytes"],
                            "kernel": "eval_bwd_kernel<2,1,7,FAST> at C3, 1 GPU",
                            "source": "profiles/" + os.path.basename(path)}
        except Exception:
            continue
    return None


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=os.environ.get("PIGS_BENCH_WORKLOAD", "C3"), choices=list(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
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


def cpu_baseline(workload, budget_s=12.0):
    """Reference's pure-PyTorch dense CPU path (port) on a bounded dense slice of the workload."""
    import torch

    from oracle import dense_torch as dt
    from tests.fixtures import synthetic
    P, N, kw, _ = WORKLOADS[workload]
    n, m = min(P, 1000), min(N, 10000)  # dense intermediates are O(n*m*d*d*4 B)
    rec = synthetic(P, min(N, 200_000), seed=0, **kw)
    rng = np.random.default_rng(7)
    gi = np.sort(rng.choice(P, n, replace=False))
    si = np.sort(rng.choice(rec["samples"].shape[0], m, replace=False))
    means, conics, values = (torch.from_numpy(rec[k][gi]) for k in ("means", "conics", "values"))
    samples = torch.from_numpy(rec["samples"][si])
    t1, pairs = dt.timed_steps(means, conics, values, samples, 1, 1)
    steps = int(max(1, min(8, budget_s / max(t1[0], 1e-3) - 2)))
    times, pairs = dt.timed_steps(means, conics, values, samples, steps, 0)
    best = min(times + t1)
    return {"value": pairs / best, "unit": UNIT, "cores": torch.get_num_threads(), "host_cpus": os.cpu_count(),
            "kind": "port",
            "sample": "dense all-pairs slice of %s: %d Gaussians x %d points, u + grad u + Hessian forward and "
                      "autograd backward, oracle/dense_torch.py restating gaussians.py:48-116; best of %d" %
                      (workload, n, m, len(times) + 1)}


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
    from pigs_b200 import _C, _lib
    from pigs_b200.sharded import ShardedGaussianSampler, allreduce_bounds, allreduce_grads
    L = _lib.lib()
    P, N, _, desc = WORKLOADS[args.workload]
    rec, ups = make_inputs(args.workload, rank, world)
    host = {k: torch.from_numpy(v).pin_memory() for k, v in rec.items()}
    host_ups = [torch.from_numpy(u).pin_memory() for u in ups]
    dev = {k: v.cuda(non_blocking=True) for k, v in host.items()}
    dev_ups = [u.cuda(non_blocking=True) for u in host_ups]
    torch.cuda.synchronize()
    D, Nl = dev["means"].shape[1], dev["samples"].shape[0]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- device-resident step (value) ----------------
    state = {}

    def step_device():
        mm = _C.sample_bounds(dev["samples"], D)
        if world > 1:
            allreduce_bounds(mm, D)
        nr, bb, sbb, rg, srg, radii = _C.preprocess_gaussians(dev["means"], dev["values"], dev["covariances"],
                                                              dev["conics"], dev["samples"], False, bounds=mm)
        a = (dev["means"], dev["values"], dev["conics"], dev["samples"], nr)
        outs = _C.sample_forward(7, *a, bb, sbb, rg, srg, False)
        dm, dv, dc = _C.sample_backward(7, *a, [dev_ups[0], dev_ups[1], dev_ups[2], None], bb, sbb, rg, srg, False)
        if world > 1:
            allreduce_grads(_C.packed_base(dm))
        state.update(rg=rg, srg=srg, outs=outs, grads=(dm, dv, dc))

    def timed(fn, steps, warmup, sampler=None):
        for _ in range(warmup):
            fn()
        barrier()
        n0 = L.pigs_launch_count()
        if sampler:
            sampler.start()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(steps):
            fn()
        b.record()
        barrier()
        if sampler:
            sampler.stop()
        ms = torch.tensor([a.elapsed_time(b)], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), L.pigs_launch_count() - n0

    clocks = ClockSampler(local)
    peaks = microbench_peaks(L, torch) if rank == 0 else None
    step_device()
    # pair count of this rank's shard from the ranges arrays, summed over ranks
    b = state["rg"].numel() // 8 - 1
    g = state["rg"].view(torch.int32)[:2 * b].view(b, 2).to(torch.int64)
    s = state["srg"].view(torch.int32)[:2 * b].view(b, 2).to(torch.int64)
    pairs_local = ((g[:, 1] - g[:, 0]) * (s[:, 1] - s[:, 0])).sum()
    pairs_t = pairs_local.clone()
    if world > 1:
        dist.all_reduce(pairs_t)
    pairs, pairs_local = int(pairs_t.item()), int(pairs_local.item())

    L.pigs_set_kernel_timing(1)
    for w in (0, 1):  # drop events recorded so far
        L.pigs_kernel_timing_read(w, ctypes.byref(ctypes.c_float()), ctypes.byref(ctypes.c_int()))
    L.pigs_set_kernel_timing(0)
    for _ in range(args.warmup):
        step_device()
    L.pigs_set_kernel_timing(1)
    total_ms, launches = timed(step_device, args.steps, 0, clocks)
    L.pigs_set_kernel_timing(0)
    kt = {}
    for w, nm in ((0, "fwd"), (1, "bwd")):
        ms, n = ctypes.c_float(), ctypes.c_int()
        _lib.check(L.pigs_kernel_timing_read(w, ctypes.byref(ms), ctypes.byref(n)), "pigs_kernel_timing_read")
        kt[nm] = (ms.value / max(n.value, 1), n.value)
    ms_per_step = total_ms / args.steps
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

    # ---------------- end-to-end step through the public API with host buffers ----------------
    e2e = None
    if not args.no_e2e:
        sampler = ShardedGaussianSampler(False, group=None) if world > 1 else None
        if sampler is None:
            from diff_gaussian_sampling import GaussianSampler
            sampler = GaussianSampler(False)
        K = D * (D + 1) // 2
        hg = torch.empty(P * (D + 1 + K), dtype=torch.float32).pin_memory()
        hu = torch.empty((Nl, 1), dtype=torch.float32).pin_memory()
        h2d = sum(host[k].numel() * 4 for k in ("means", "values", "covariances", "conics", "samples")) + \
            sum(u.numel() * 4 for u in host_ups)
        d2h = hg.numel() * 4 + hu.numel() * 4

        def step_e2e():
            d = {k: host[k].cuda(non_blocking=True) for k in ("means", "values", "covariances", "conics", "samples")}
            gu = [u.cuda(non_blocking=True) for u in host_ups]
            for k in ("means", "values", "conics"):
                d[k].requires_grad_(True)
            sampler.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
            u = sampler.sample_gaussians()
            ux = sampler.sample_gaussians_derivative()
            uxx = sampler.sample_gaussians_laplacian()
            loss = (u * gu[0]).sum() + (ux * gu[1]).sum() + (uxx * gu[2]).sum()
            loss.backward()
            hg[:P * D].copy_(d["means"].grad.reshape(-1), non_blocking=True)
            hg[P * D:P * (D + 1)].copy_(d["values"].grad.reshape(-1), non_blocking=True)
            hg[P * (D + 1):].copy_(d["conics"].grad.reshape(-1), non_blocking=True)
            hu.copy_(u.detach(), non_blocking=True)
            torch.cuda.current_stream().synchronize()

        e_ms, _ = timed(step_e2e, args.steps, max(args.warmup, 2))
        e2e = {"value": pairs / (e_ms / args.steps * 1e-3), "unit": UNIT, "ms_per_step": e_ms / args.steps,
               "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
               "api": "diff_gaussian_sampling.GaussianSampler.preprocess/sample_gaussians/_derivative/_laplacian + loss.backward()"}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    ck = clocks.summary()
    fmax = (ck.get("sm_max_mhz") or 1965.0) * 1e6
    peak = sms * 128 * fmax / 1e12
    bwd_ms, fwd_ms = kt["bwd"][0], kt["fwd"][0]
    ach_bwd = pairs_local * LANE_OPS_BWD / (bwd_ms * 1e-3) / 1e12 if bwd_ms > 0 else None
    ach_all = pairs_local * (LANE_OPS_FWD + LANE_OPS_BWD) / ((bwd_ms + fwd_ms) * 1e-3) / 1e12 if bwd_ms > 0 else None
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic (seeded jittered-grid Gaussians, uniform points; tests/fixtures.synthetic seed 0)",
        "config": {"workload": "%s: %s" % (args.workload, desc), "gaussians": P, "points": N, "points_per_gpu": Nl,
                   "pairs": pairs, "functions": "u + grad u + Hessian fused fwd, fused bwd to means/conics/values",
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
                     "traffic": (ncu_traffic() or {}).get("dram_bytes_per_launch"),
                     "traffic_note": (ncu_traffic() or {}).get("kernel", "no committed ncu summary") + "; " +
                                     (ncu_traffic() or {}).get("source", "")},
        "clocks": ck,
    }
    if D != 2:
        # the lane-op counts above are the 2-D one-channel kernels'; the 1-D case (C1) is launch-bound at its size
        out["roofline"] = {"bound": "launch latency", "note": "1-D, 1K x 10K: ~60 dependent launches of a few us each; "
                           "no FP32 roofline is claimed for this workload", "kernel_ms": {"eval_fwd": fwd_ms, "eval_bwd": bwd_ms}}
    if e2e:
        out["e2e"] = e2e
    if world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(args.workload)
    emit(out)
    if world > 1:
        dist.destroy_process_group()


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path — its pure-PyTorch dense sampling
    (gaussians.py:48-116; the module itself cannot travel to the GPU box, so oracle/dense_torch.py restates it) —
    on the host cores with all the threads torch uses, on the same workload's data.  Each step is a bounded
    sample: a dense 1000-Gaussian x 10000-point slice (dense intermediates are O(n*m*d^2))."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    import torch

    from oracle import dense_torch as dt
    from tests.fixtures import synthetic
    P, N, kw, desc = WORKLOADS[args.workload]
    n, m = min(P, 1000), min(N, 10000)
    rec = synthetic(P, min(N, 200_000), seed=0, **kw)
    rng = np.random.default_rng(7)
    gi = np.sort(rng.choice(P, n, replace=False))
    si = np.sort(rng.choice(rec["samples"].shape[0], m, replace=False))
    means, conics, values = (torch.from_numpy(rec[k][gi]) for k in ("means", "conics", "values"))
    samples = torch.from_numpy(rec["samples"][si])
    steps, warmup = min(args.steps, 20), min(args.warmup, 3)
    t0 = time.perf_counter()
    times, pairs = dt.timed_steps(means, conics, values, samples, steps, warmup)
    ms = 1e3 * sum(times) / len(times)
    value = pairs / (ms * 1e-3)
    sample = ("each step = dense all-pairs slice of %s: %d Gaussians x %d points, u + grad u + Hessian forward + autograd "
              "backward (pure-PyTorch path of gaussians.py:48-116, restated in oracle/dense_torch.py)" % (args.workload, n, m))
    out = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps,
           "warmup": warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
           "dtype": "f32", "data": "synthetic (same seeded recipe as the GPU arm)",
           "config": {"workload": "%s: %s" % (args.workload, desc), "gaussians": P, "points": N, "sample": sample},
           "cpu_baseline": {"value": value, "unit": UNIT, "cores": torch.get_num_threads(), "host_cpus": os.cpu_count(),
                            "kind": "port", "sample": sample},
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
