"""Dev tool: split a kernel's SASS (ncu --page source --csv export) into contiguous regions of similar execution count
and print each region's share of executed warp instructions and of stall samples.
usage: ncu -i rep.ncu-rep --page source --csv --kernel-name regex:NAME > src.csv; python tools/ncu_source_regions.py src.csv"""
import csv
import math
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = next(r for r in rows if "Address" in r and "Source" in r)
isrc, ie, iss = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Warp Stall Sampling (All Samples)")
data = []
for r in rows:
    if len(r) == len(hdr) and r[ie].isdigit():
        data.append(r)
    elif data and len(r) == len(hdr):
        break  # second view of the same kernel
tot = sum(int(r[ie]) for r in data)
tots = sum(int(r[iss]) for r in data)
print("instructions", len(data), "executed", tot, "samples", tots)
seg, cur = [], None
for i, r in enumerate(data):
    e = int(r[ie])
    lvl = 0 if e == 0 else round(3 * math.log10(e))
    if cur is None or cur[0] != lvl:
        if cur:
            seg.append(cur)
        cur = [lvl, i, i, e, int(r[iss])]
    else:
        cur[2] = i
        cur[3] += e
        cur[4] += int(r[iss])
seg.append(cur)
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.01
for s in seg:
    if s[3] > tot * thr or s[4] > tots * thr:
        print("instr %4d-%4d  exec %5.1f%%  samples %5.1f%%  n=%-4d first: %s" % (
            s[1], s[2], 100 * s[3] / tot, 100 * s[4] / max(tots, 1), s[2] - s[1] + 1, data[s[1]][isrc].strip()[:70]))
