#!/usr/bin/env python3
"""Per-source-line executed-instruction histogram from `ncu --page source --csv --print-source sass,cuda`."""
import csv
import sys
from collections import defaultdict

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur_file = None
hdr = None
agg = defaultdict(lambda: [0, 0, "", 0])
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        hdr = None
        continue
    if r[0] == "Function Name":
        continue
    if r[0] == "Line No":
        hdr = r
        ie = hdr.index("Instructions Executed")
        te = hdr.index("Thread Instructions Executed")
        sm = hdr.index("# Samples")
        continue
    if hdr is None or len(r) <= ie or not r[0].isdigit():      # skip the SASS rows listed under each source line
        continue
    try:
        n = int(r[ie])
        s = int(r[sm])
    except ValueError:
        continue
    key = (cur_file, int(r[0]) if r[0].isdigit() else -1)
    agg[key][0] += n
    agg[key][1] += s
    try:
        agg[key][3] += int(r[te])
    except ValueError:
        pass
    if r[1]:
        agg[key][2] = r[1]
tot = sum(v[0] for v in agg.values())
tots = sum(v[1] for v in agg.values())
print("total warp instructions", tot, "samples", tots)
byfile = defaultdict(int)
for (f, l), v in agg.items():
    byfile[f] += v[0]
print(dict(byfile))
for (f, l), v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{v[0]:>11} {100 * v[0] / tot:5.1f}%  smp {100 * v[1] / max(tots, 1):5.1f}%  thr {v[3] / max(v[0], 1):4.1f}  {f}:{l}: {v[2].strip()[:90]}")
