"""Dev tool: static issue-cost model of a kernel's hot loop from its SASS.

Measured on B200 (tools/mb_issue.cu): an FP32 instruction whose source operands need THREE register-file reads
(three distinct registers, none served by the operand reuse cache) occupies the dispatch port for 2 cycles;
instructions with <= 2 new register reads (immediates, constants, uniform registers and .reuse hits are free)
issue at 1/clk; MUFU.EX2 costs ~1.25.  The model reproduces the measured issue rate of the evaluation kernels
to within a few percent, so loop-order experiments can be scored here without a GPU.

usage: sass_issue_model.py <object-or-so> <substring of mangled kernel name> [-v]
"""
import re
import subprocess
import sys


def kernel_sass(path, pat):
    out = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True).stdout
    cur, keep, name = [], None, None
    for line in out.splitlines():
        if "Function :" in line:
            if keep is not None and cur:
                break
            name = line.split("Function :")[1].strip()
            keep = True if pat in name else None
            cur = []
        elif keep:
            m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", line)
            if m:
                cur.append((int(m.group(1), 16), m.group(2).strip()))
    return name, cur


def srcs(t):
    t = re.sub(r"^@!?U?P\d+\s+", "", t)
    op = t.split()[0]
    ops = t[len(op):].split(",")
    out = []
    for o in ops[1:]:
        m = re.search(r"(?<![U])R(\d+)(\.reuse)?", o)
        # on packed FP32x2 instructions a `.F32` operand is a scalar broadcast: it does not count towards the
        # three-pair-read penalty (tools/mb_issue.cu: two pairs + one broadcast = 2 cycles, three pairs = 3)
        bcast = op.endswith("2") and ".F32x2" not in o and ".F32" in o
        out.append((int(m.group(1)), bool(m.group(2))) if (m and not bcast) else None)
    return out


def hot_loop(ins):
    best = None
    for i, (a, t) in enumerate(ins):
        m = re.search(r"BRA(?:\.U)?\s+(?:!?U?P\d+,\s*)?(0x[0-9a-f]+)", t)
        if not m:
            continue
        tgt = int(m.group(1), 16)
        if tgt >= a:
            continue
        body = [(x, y) for x, y in ins if tgt <= x <= a]
        nm = sum(1 for _, y in body if y.startswith("MUFU.EX2"))
        nf = sum(1 for _, y in body if y.startswith(("FFMA", "FMUL", "FADD")))
        if nm >= 1 and nf >= 8 * nm:
            if best is None or len(body) < len(best):
                best = body
    return best


def model(body, verbose=False, persistent=False):
    """persistent=False: a reuse-cache entry serves only the immediately following instruction (pessimistic);
    persistent=True: it stays until another .reuse operand replaces it in that slot or its register is
    overwritten (fits the measured kernels slightly better: never above the measured cycles)."""
    cache = [None] * 8
    extra = 0
    ops = {}
    for _, t in body:
        s = srcs(t)
        need = []
        for slot, x in enumerate(s):
            if x is None:
                continue
            if cache[slot] == x[0]:
                continue
            need.append(x[0])
        need = list(dict.fromkeys(need))
        c = len(need) >= 3
        extra += c
        op = re.sub(r"^@!?U?P\d+\s+", "", t).split()[0]
        ops[op] = ops.get(op, 0) + 1
        if verbose:
            print(("** " if c else "   ") + t)
        if persistent:
            for slot, x in enumerate(s):
                if x and x[1]:
                    cache[slot] = x[0]
            tt = re.sub(r"^@!?U?P\d+\s+", "", t)
            dm = re.search(r"(?<![U])R(\d+)", tt[len(tt.split()[0]):].split(",")[0])
            if dm:
                cache = [None if c == int(dm.group(1)) else c for c in cache]
        else:
            nc = [None] * 8
            for slot, x in enumerate(s):
                if x and x[1]:
                    nc[slot] = x[0]
            cache = nc
    return extra, ops


if __name__ == "__main__":
    name, ins = kernel_sass(sys.argv[1], sys.argv[2])
    body = hot_loop(ins)
    extra, ops = model(body, "-v" in sys.argv)
    extra_p, _ = model(body, "-vp" in sys.argv, True)
    n = len(body)
    mufu = ops.get("MUFU.EX2", 0)
    fp = sum(v for k, v in ops.items() if k.startswith(("FFMA", "FMUL", "FADD")))
    packed = sum(v for k, v in ops.items() if k in ("FFMA2", "FMUL2", "FADD2"))
    n_issue = n + packed  # a packed FP32x2 instruction occupies the dispatch port for two cycles
    lane_ops = fp + packed
    cyc = n_issue + extra + 0.25 * mufu
    print(name)
    print("loop: %d instr (%d FP32, %d MUFU.EX2, %d other), +%d three-read stalls -> %.1f cycles; per MUFU (pair): "
          "%.2f instr, %.2f cycles; FP32 lane-ops per cycle %.3f" % (n, fp, mufu, n - fp - mufu, extra, cyc, n / mufu, cyc / mufu, lane_ops / cyc))
    print("  of which packed FP32x2: %d (2 cycles each, 2 lane-ops each); FP32 lane-ops per pair %.2f" % (packed, lane_ops / mufu))
    cyc_p = n_issue + extra_p + 0.25 * mufu
    print("persistent-reuse variant: +%d stalls -> %.2f cycles per pair, FP32 lane-ops per cycle %.3f" % (extra_p, cyc_p / mufu, lane_ops / cyc_p))
