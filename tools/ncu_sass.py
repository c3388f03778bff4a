#!/usr/bin/env python3
"""Linear annotated SASS of the hot code: ncu_sass.py src.csv [min_count] -> address order, with source file:line and executions."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
thr = int(sys.argv[2]) if len(sys.argv) > 2 else 100000
cur = None; line = None; hdr = None
out = {}
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur = r[1].split('/')[-1]; hdr = None; continue
    if r[0] == "Line No": hdr = r; ie = hdr.index("Instructions Executed"); continue
    if hdr is None: continue
    if r[0].isdigit(): line = int(r[0]); continue
    if len(r) > ie and r[2].startswith("0x"):
        try: n = int(r[ie])
        except ValueError: continue
        a = int(r[2], 16)
        if a not in out: out[a] = (r[3].strip(), n, cur, line)
base = min(out) if out else 0
tot = 0
for a in sorted(out):
    s, n, f, l = out[a]
    if n >= thr:
        tot += n
        print(f"{a-base:6x} {n:>10d}  {f}:{l:<4d} {s}")
print("sum", tot)
