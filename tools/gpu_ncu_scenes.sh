#!/bin/bash
# short-metric ncu captures of the steady-state (cost-ordered) launch on the other config scenes
mkdir -p gpurun_out
METRICS=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct,dram__bytes_read.sum,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio,smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio
for s in animemaid valorant_ghost dense_monkey; do
  timeout 900 ncu --metrics $METRICS --clock-control none -k regex:k_render --launch-skip 4 -c 1 -o gpurun_out/ncu_$s -f python tools/sweep.py --scenes $s --spp 64 --variants -1 --blocks 128 --reps 3 > gpurun_out/ncu_$s.log 2>&1
done
ls -la gpurun_out/ncu_*.ncu-rep
