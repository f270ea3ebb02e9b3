#!/usr/bin/env python
"""Top instructions by warp-stall samples from `ncu -i X.ncu-rep --page source --csv` (SASS view)."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]
ci, si, ei = h.index("Warp Stall Sampling (All Samples)"), h.index("Source"), h.index("Instructions Executed")
stall_cols = [i for i, n in enumerate(h) if n.startswith("stall_") and "Not Issued" not in n]
data = []
for k, r in enumerate(rows[hi + 1:]):
    try:
        data.append((float(r[ci]), k, r))
    except (ValueError, IndexError):
        pass
tot = sum(d[0] for d in data) or 1
print("total samples", tot, "instructions", len(data))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
for v, k, r in sorted(data, key=lambda x: -x[0])[:top]:
    reasons = sorted(((float(r[i] or 0), h[i][6:]) for i in stall_cols), reverse=True)[:2]
    print(f"{v:8.0f} {100 * v / tot:5.1f}% #{k:4d} exec {r[ei]:>8s} {r[si].strip()[:70]:70s} {reasons}")
