"""Dev tool: condense an `ncu --set full` report into the handful of counters DESIGN.md argues from.
usage: python tools/ncu_summary.py <report.ncu-rep> <out.json>   (needs ncu on PATH; no GPU)"""
import csv
import io
import json
import subprocess
import sys

KEYS = {
    "gpu__time_duration.sum": "duration",
    "launch__registers_per_thread": "registers_per_thread",
    "launch__grid_size": "grid",
    "launch__block_size": "block",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "achieved_occupancy_pct",
    "smsp__issue_active.avg.per_cycle_active": "issue_slots_busy_per_cycle",
    "smsp__warps_eligible.avg.per_cycle_active": "eligible_warps_per_cycle",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active": "pipe_fma_pct",
    "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed": "pipe_fmaheavy_cycles_pct",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active": "pipe_xu_mufu_pct",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active": "pipe_alu_pct",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active": "pipe_lsu_pct",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed": "shared_mem_wavefronts_pct",
    "smsp__inst_executed.sum": "warp_instructions",
    "dram__bytes_read.sum": "dram_read",
    "dram__bytes_write.sum": "dram_write",
    "lts__t_bytes.sum": "l2_bytes",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_throughput_pct",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram_throughput_pct",
    "smsp__cycles_active.avg": "smsp_cycles_active",
    "l1tex__t_sector_hit_rate.pct": "l1_hit_rate_pct",
    "lts__t_sector_hit_rate.pct": "l2_hit_rate_pct",
    "lts__t_sectors.sum": "l2_sectors",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum": "l1_global_load_sectors",
    "launch__occupancy_limit_registers": "occupancy_limit_registers",
    "smsp__inst_executed_per_warp.ratio": "inst_per_warp",
}
STALLS = "smsp__pcsamp_warps_issue_stalled_"


def to_bytes(v, unit):
    v = float(v.replace(",", ""))
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


def main():
    rep, out = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    res = []
    for r in rows[2:]:
        k = {"kernel": r[idx["Kernel Name"]].split("(")[0]}
        for m, name in KEYS.items():
            if m in idx:
                u = units[idx[m]]
                if "byte" in u:
                    k[name + "_bytes"] = to_bytes(r[idx[m]], u)
                else:
                    try:
                        k[name] = float(r[idx[m]].replace(",", ""))
                    except ValueError:
                        k[name] = r[idx[m]]
                    if name == "duration":
                        k["duration_unit"] = u
        st = {}
        for h, i in idx.items():
            if h.startswith(STALLS) and "not_issued" not in h:
                try:
                    st[h[len(STALLS):]] = int(float(r[i].replace(",", "")))
                except ValueError:
                    pass
        tot = sum(st.values()) or 1
        k["warp_state_samples_pct"] = {a: round(100.0 * b / tot, 1) for a, b in sorted(st.items(), key=lambda x: -x[1])[:8]}
        res.append(k)
    import hashlib
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hh = hashlib.sha1()
    for name in ("eval_bwd.cu", "eval_fwd.cu", "eval_common.cuh", "common.cuh", "pack.cu"):  # = bench.py EVAL_SOURCES
        with open(os.path.join(root, "pigs_b200", "csrc", name), "rb") as f:
            hh.update(f.read())
    json.dump({"report": rep, "source_sha": hh.hexdigest()[:16], "note": "ncu --set full --clock-control none; per-launch values; durations under the "
               "profiler are serialised replays, not bench numbers", "kernels": res}, open(out, "w"), indent=1)
    for k in res:
        print(k["kernel"][:60], k.get("duration"), k.get("duration_unit"), "issue", k.get("issue_slots_busy_per_cycle"),
              "fma%", k.get("pipe_fma_pct"), "dramR MB", round(k.get("dram_read_bytes", 0) / 1e6, 1), "dramW MB",
              round(k.get("dram_write_bytes", 0) / 1e6, 1))


if __name__ == "__main__":
    main()
