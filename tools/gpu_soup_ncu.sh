#!/bin/bash
# ncu --set full capture of the render kernel on soup-10M (config 5, the HBM-bound regime), 4 spp at 3840x2160.
set -x
mkdir -p gpurun_out
V=${SOUP_VARIANT:-7}
timeout 900 python tools/run_configs.py --configs 5 --spp-scale 0.0625 --variant $V > gpurun_out/soup_plain.jsonl 2> gpurun_out/soup_plain.err &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:k_render --launch-skip 1 -c 2 -o gpurun_out/soup_full -f \
  python tools/run_configs.py --configs 5 --spp-scale 0.0625 --variant $V > gpurun_out/soup_ncu.log 2>&1
