#!/bin/bash
# On the GPU box: time and count warp instructions of every ladder build (tools/build_ladder.sh) at n sessions.
n=${1:-10000}
for r in 0 1 2 3 4 5 6; do
  export DVB_LIB=$PWD/_ab/rung$r.so
  ms=$(timeout 120 python tools/prof_bind.py $n 4 2>&1 | grep "kernel ms" | tail -2 | awk '{s+=$3} END {print s/NR}')
  timeout 300 ncu --metrics smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__thread_inst_executed_per_inst_executed.ratio --clock-control none -k regex:k_bind_facts -s 1 -c 1 --csv --log-file gpurun_out/ladder_$r.csv python tools/prof_bind.py $n 2 > /dev/null 2>&1
  python - "$r" "$ms" gpurun_out/ladder_$r.csv <<'PY'
import csv, sys, json
r, ms, path = sys.argv[1], float(sys.argv[2]), sys.argv[3]
rows = [x for x in csv.reader(l for l in open(path) if not l.startswith("==")) if x]
hdr = rows[0]; m = {x[hdr.index("Metric Name")]: float(x[hdr.index("Metric Value")].replace(",", "")) for x in rows[1:]}
print(json.dumps({"rung": int(r), "kernel_ms": round(ms, 4), "warp_inst": int(m["smsp__inst_executed.sum"]),
                  "issue_active_pct": m["smsp__issue_active.avg.pct_of_peak_sustained_active"],
                  "lanes_per_inst": m["smsp__thread_inst_executed_per_inst_executed.ratio"]}))
PY
done
