#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel name (share of the captured time)."""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
hdr = rows[hdr_i]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[hdr_i + 1:]:
    if len(r) < len(hdr):
        continue
    v = float(r[vi].replace(",", ""))
    scale = {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3}.get(r[ui], 1e-3)
    name = re.sub(r"\(.*", "", r[ki])
    agg[name][0] += 1
    agg[name][1] += v * scale
tot = sum(v[1] for v in agg.values())
print(f"{'kernel':70s} {'n':>5s} {'total ms':>10s} {'avg us':>9s} {'share':>6s}")
for n, (c, v) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{n[:70]:70s} {c:5d} {v / 1e3:10.3f} {v / c:9.1f} {100 * v / tot:5.1f}%")
print(f"total {tot / 1e3:.3f} ms over {sum(v[0] for v in agg.values())} launches")
