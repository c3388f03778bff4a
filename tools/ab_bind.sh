#!/bin/bash
# A/B of libdvb200 builds on the bind kernel: time (plain run) then instruction count (ncu metrics run) per library.
# usage: tools/ab_bind.sh n_sessions lib1 [lib2 ...]   ("default" = the in-tree library)
n=$1; shift
for lib in "$@"; do
  if [ "$lib" = default ]; then unset DVB_LIB; else export DVB_LIB=$PWD/$lib; fi
  echo "== $lib"
  timeout 120 python tools/prof_bind.py $n 3 2>&1 | tail -3
  timeout 300 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,launch__registers_per_thread,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:k_bind_ -s 2 -c 2 --csv --log-file gpurun_out/ab.csv python tools/prof_bind.py $n 2 > /dev/null 2>&1
  grep -v "^==" gpurun_out/ab.csv | python -c "
import csv,sys
for r in list(csv.reader(sys.stdin))[1:]: print('   ', r[-3], r[-1])"
done
