#!/bin/bash
# A/B of library builds under ab/*.so on the BASELINE configs (same box, alternating). Usage: tools/ab.sh "1,2,3" 0.25 [rounds]
CFG=${1:-1,2,3,4a,4b}; SCALE=${2:-0.25}; ROUNDS=${3:-2}
cp dustraytracer_b200/libdrt.so /tmp/libdrt_keep.so
for r in $(seq 1 $ROUNDS); do
  for so in ab/*.so; do
    name=$(basename $so .so)
    cp $so dustraytracer_b200/libdrt.so
    timeout 900 python tools/run_configs.py --configs $CFG --spp-scale $SCALE --ref-frames 0 2>/dev/null | grep '^{' | sed "s/^{/{\"lib\": \"$name\", \"round\": $r, /" >> gpurun_out/ab.jsonl
  done
done
cp /tmp/libdrt_keep.so dustraytracer_b200/libdrt.so
