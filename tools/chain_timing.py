"""Per-stage device time of the preprocess chain (pigs_set_stage_timing): what the tile binning of one workload really
spends where, kernels back to back as in production (an ncu launch list times every kernel alone with cold caches).
python tools/chain_timing.py [C3|C2|C1|shard8] > profiles/rNN_chain_timing.json"""
import ctypes
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pigs_b200 import GaussianSampler, _lib  # noqa: E402
from tests.fixtures import synthetic  # noqa: E402

CASES = {"C3": (1_000_000, 4_000_000, {}), "C2": (100_000, 1_000_000, {}), "C1": (1000, 10000, dict(D=1)),
         "shard8": (1_000_000, 125_000, {})}


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "C3"
    P, N, kw = CASES[name]
    rec = synthetic(P, N, seed=0, **kw)
    d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
    s = GaussianSampler(False)
    L = _lib.lib()
    flush = torch.empty(64 * 1024 * 1024, dtype=torch.float32, device="cuda")
    for _ in range(3):
        s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
    acc, totals = {}, []
    names = (ctypes.c_char_p * 64)()
    ms = (ctypes.c_float * 64)()
    for it in range(9):
        flush.fill_(1.0)
        torch.cuda.synchronize()
        L.pigs_set_stage_timing(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
        b.record()
        torch.cuda.synchronize()
        n = L.pigs_stage_timing_read(names, ms, 64)
        L.pigs_set_stage_timing(0)
        totals.append(a.elapsed_time(b))
        seen = {}
        for i in range(1, n):
            nm = names[i].decode()
            k = seen.get(nm, 0)
            seen[nm] = k + 1
            acc.setdefault("%s #%d" % (nm, k + 1) if nm.startswith("radix") else nm, []).append(ms[i])
    rows = [{"stage": k, "us": round(1e3 * float(np.median(v)), 2)} for k, v in acc.items()]
    print(json.dumps({"gpu": torch.cuda.get_device_name(0), "workload": name, "P": P, "N": N, "R": s.num_rendered,
                      "preprocess_total_us": round(1e3 * float(np.median(totals)), 1),
                      "sum_of_stages_us": round(sum(r["us"] for r in rows), 1),
                      "timing": "cudaEvents recorded by the library behind each stage, kernels back to back on the stream, L2 "
                                "flushed before each of 9 repetitions, medians",
                      "stages": rows}, indent=1))


if __name__ == "__main__":
    main()
