#!/bin/bash
# per-library ncu counters of k_bind_claims: instructions, issue rate, stall reasons per issue.  usage: tools/ab_stalls.sh n lib...
n=$1; shift
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__thread_inst_executed_per_inst_executed.ratio,sm__warps_active.avg.pct_of_peak_sustained_active
for s in barrier branch_resolving long_scoreboard math_pipe_throttle no_instruction not_selected short_scoreboard wait dispatch_stall mio_throttle lg_throttle; do M=$M,smsp__average_warps_issue_stalled_${s}_per_issue_active.ratio; done
for lib in "$@"; do
  if [ "$lib" = default ]; then unset DVB_LIB; else export DVB_LIB=$PWD/$lib; fi
  timeout 300 ncu --metrics $M --clock-control none -k regex:k_bind_ -s 2 -c 2 --csv --log-file gpurun_out/st.csv python tools/prof_bind.py $n 2 > /dev/null 2>&1
  echo "== $lib"
  python tools/ncu_metrics_table.py gpurun_out/st.csv
done
