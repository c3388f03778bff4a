#!/bin/bash
# no_instruction / issue rate of the other kernels (gate, index, annotation): one ncu metrics run per tool
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__thread_inst_executed_per_inst_executed.ratio,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread
for s in no_instruction wait long_scoreboard short_scoreboard barrier math_pipe_throttle branch_resolving; do M=$M,smsp__average_warps_issue_stalled_${s}_per_issue_active.ratio; done
run() { # name regex cmd...
  local name=$1 rx=$2; shift; shift
  timeout 300 ncu --metrics $M --clock-control none -k regex:$rx -c 6 --csv --log-file gpurun_out/so_$name.csv "$@" > /dev/null 2>&1
  echo "== $name"; python tools/ncu_metrics_table.py gpurun_out/so_$name.csv | grep -E "time|no_inst|issue_active|inst_executed.sum|registers|per_inst"
}
run gate "k_question_keys|k_gate_match" python tools/bench_gate.py --candidates 200000 --history 50000 --reps 1
run index "k_ix_score|k_symbolize|k_rs_scatter" python tools/bench_index.py --chunks 200000 --queries 2000 --reps 1
run annot "k_text_features|k_annotate_turns|k_contains" python tools/bench_annot.py
