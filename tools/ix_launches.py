#!/usr/bin/env python3
"""Per-kernel table of an `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv` launch list."""
import csv, sys
from collections import defaultdict
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if not l.startswith('=='))]
h = rows[0]; ki = h.index('Kernel Name'); mi = h.index('Metric Name'); vi = h.index('Metric Value'); ii = h.index('ID')
d = defaultdict(dict)
for r in rows[1:]:
    d[(int(r[ii]), r[ki])][r[mi]] = float(r[vi].replace(',', ''))
tot = 0
for (i, k), m in sorted(d.items()):
    t = m.get('gpu__time_duration.sum', 0); tot += t
    print(f"{i:3d} {k[:40]:40s} {t/1e3:9.1f} us  rd {m.get('dram__bytes_read.sum',0)/1e6:9.1f} MB  wr {m.get('dram__bytes_write.sum',0)/1e6:9.1f} MB")
print(f"total {tot/1e6:.3f} ms")
