#!/usr/bin/env python
"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv): per kernel count, total, share, mean."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1], errors="replace")))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
h = rows[hdr]
ki, vi, gi, bi = h.index("Kernel Name"), h.index("Metric Value"), h.index("Grid Size"), h.index("Block Size")
ui = h.index("Metric Unit")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[hdr + 1:]:
    if len(r) <= vi or not r[vi]:
        continue
    scale = {"ns": 1.0, "us": 1e3, "usecond": 1e3, "nsecond": 1.0, "ms": 1e6, "msecond": 1e6}.get(r[ui], 1.0)
    k = (r[ki][:90], r[gi], r[bi])
    agg[k][0] += 1
    agg[k][1] += float(r[vi].replace(",", "")) * scale
tot = sum(v[1] for v in agg.values())
print("%-92s %-16s %-14s %6s %10s %6s %9s" % ("kernel", "grid", "block", "n", "total_us", "share", "mean_us"))
for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
    print("%-92s %-16s %-14s %6d %10.1f %5.1f%% %9.2f" % (k[0], k[1], k[2], v[0], v[1] / 1e3, 100 * v[1] / tot, v[1] / v[0] / 1e3))
print("total %.1f us over %d launches" % (tot / 1e3, sum(v[0] for v in agg.values())))
