#!/usr/bin/env python
"""Group the SASS of an ncu source-page CSV into runs of equal execution count and sum their stall samples."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]; ci, si, ei = h.index("Warp Stall Sampling (All Samples)"), h.index("Source"), h.index("Instructions Executed")
data = [(float(r[ci] or 0), int(r[ei] or 0), r[si].strip()) for r in rows[hi+1:] if len(r) > ci]
start = 0
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 30
for k in range(1, len(data)+1):
    if k == len(data) or data[k][1] != data[start][1]:
        s = sum(d[0] for d in data[start:k])
        if s >= thr: print(f"[{start:5d},{k:5d}) n={k-start:4d} exec={data[start][1]:9d} samples={s:7.0f}  {data[start][2][:60]}")
        start = k
