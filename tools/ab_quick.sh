#!/bin/bash
# usage: tools/ab_quick.sh <world> <W> <H> <pp> lib_or_env ...   -- time + instruction counts for several builds/envs
w=$1; W=$2; H=$3; pp=$4; shift 4
for spec in "$@"; do
  echo "== $spec"
  env $spec python tools/profile_render.py --world $w --width $W --height $H --pp $pp --frames 3 | tail -1
  env $spec ncu --metrics smsp__inst_executed.sum,smsp__thread_inst_executed.sum,gpu__time_duration.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio --clock-control none -k regex:k_render -s 1 -c 1 python tools/profile_render.py --world $w --width $W --height $H --pp $pp --frames 2 2>/dev/null | grep -E "inst_executed|duration|issue_active|no_instruction" | awk '{printf "   %s %s %s\n", $1, $(NF-1), $NF}'
done
