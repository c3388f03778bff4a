#!/usr/bin/env python3
"""Prints kernel / metric / value rows of an `ncu --csv --metrics ...` log: ncu_metrics_table.py log.csv"""
import csv, sys
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if not l.startswith("==")) if r]
h = rows[0]
for r in rows[1:]:
    n = r[h.index("Metric Name")].replace("smsp__average_warps_issue_stalled_", "st_").replace("_per_issue_active.ratio", "")
    print("   %-22s %-58s %s" % (r[h.index("Kernel Name")][:22], n, r[h.index("Metric Value")]))
