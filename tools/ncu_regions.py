#!/usr/bin/env python3
"""Sum executed warp instructions of an ncu source csv over line ranges: ncu_regions.py src.csv file:lo-hi[:label] ..."""
import csv, sys
from collections import defaultdict
rows = list(csv.reader(open(sys.argv[1])))
cur = None; hdr = None
agg = defaultdict(int)
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur = r[1].split('/')[-1]; hdr = None; continue
    if r[0] == "Line No": hdr = r; ie = hdr.index("Instructions Executed"); continue
    if hdr is None or not r[0].isdigit(): continue
    try: n = int(r[ie])
    except ValueError: continue
    agg[(cur, int(r[0]))] += n
tot = sum(agg.values())
print("total", tot)
for spec in sys.argv[2:]:
    parts = spec.split(":")
    f, rg = parts[0], parts[1]
    lo, hi = map(int, rg.split("-"))
    v = sum(x for (ff, l), x in agg.items() if ff == f and lo <= l <= hi)
    print(f"{spec:40s} {v:>12d} {100*v/tot:5.1f}%")
