import csv, sys
for f in sys.argv[1:]:
    rows = list(csv.reader(open(f)))
    hi = next(i for i,r in enumerate(rows) if r and r[0]=="ID")
    h = rows[hi]; ki, vi, gi = h.index("Kernel Name"), h.index("Metric Value"), h.index("Grid Size")
    data = [(r[ki][:30], r[gi], float(r[vi].replace(",",""))) for r in rows[hi+1:] if len(r) >= len(h)]
    half = len(data)//2
    print(f, "total ms %.3f" % (sum(v for _,_,v in data[half:])/1e6))
    print(" ".join(f"{g.split(',')[0][1:]}:{v/1e3:.0f}" for n,g,v in data[half:]))
