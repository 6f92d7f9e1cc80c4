"""Dev tool: condense an ncu launch list (`--metrics gpu__time_duration.sum --csv`) into per-kernel totals and shares.
usage: python tools/launch_summary.py <launches.csv> > profiles/rNN_launches_summary.txt"""
import csv
import re
import sys
from collections import OrderedDict

rows = [r for r in csv.reader(open(sys.argv[1])) if r and r[0] != "==PROF=="]
start = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
hdr = rows[start]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
acc = OrderedDict()
for r in rows[start + 1:]:
    if len(r) <= vi:
        continue
    name = re.sub(r"\(.*", "", r[ki])[:100]
    v = float(r[vi].replace(",", ""))
    v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[ui], 1e-6)
    n, t = acc.get(name, (0, 0.0))
    acc[name] = (n + 1, t + v)
total = sum(t for _, t in acc.values())
print("# launch list under ncu (gpu__time_duration.sum, --clock-control none): per-launch times are cold-cache and")
print("# serialised; what must agree with bench.py is each kernel's SHARE of the step")
print("# launches   total_ms   avg_us   share%   kernel")
for name, (n, t) in acc.items():
    print("%6d %10.3f %10.1f %7.2f   %s" % (n, t, 1e3 * t / n, 100 * t / total, name))
