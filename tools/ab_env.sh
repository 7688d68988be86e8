#!/bin/bash
# Sweep of DRT_SCHED_BIAS values on the current library. Usage: tools/ab_env.sh "6 -1 -3" "2,3,4a" 0.25
VALS=${1:-"6"}; CFG=${2:-2,3,4a,4b,5s}; SCALE=${3:-0.25}
for r in 1 2; do for v in $VALS; do
  DRT_SCHED_BIAS="$v,0,0" timeout 900 python tools/run_configs.py --configs $CFG --spp-scale $SCALE --ref-frames 0 2>/dev/null | grep '^{' | sed "s/^{/{\"bias\": $v, \"round\": $r, /" >> gpurun_out/ab_env.jsonl
done; done
