#!/usr/bin/env python
"""Per-kernel totals of an ncu launch list (`ncu --metrics gpu__time_duration.sum --csv --log-file X`):
    python tools/launch_summary.py gpurun_out/launches.csv "command line" > profiles/..._summary.txt"""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(open(sys.argv[1], errors="ignore")) if r]
hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
h = rows[hdr]
kn, mv, mu = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
tot = collections.OrderedDict()
for r in rows[hdr + 1:]:
    if len(r) <= mv:
        continue
    name = re.sub(r"\(.*$", "", r[kn]).strip()
    us = float(r[mv].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r[mu], 1.0)
    c = tot.setdefault(name, [0, 0.0])
    c[0] += 1
    c[1] += us
allus = sum(v[1] for v in tot.values())
print(sys.argv[2] if len(sys.argv) > 2 else "")
print("(cold-cache, serialised launches under ncu: compare shares, not absolutes)")
print("%-95s %6s %12s %7s %10s" % ("kernel", "count", "total_us", "share", "avg_us"))
for name, (c, us) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print("%-95s %6d %12.1f %6.1f%% %10.2f" % (name[:95], c, us, 100 * us / allus, us / c))
