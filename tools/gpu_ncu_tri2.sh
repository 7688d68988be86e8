#!/bin/bash
# ncu evidence for the two-triangle TRI step: variant 4 and variant 15 on dense_monkey.glb, 1080p, 16 spp.
mkdir -p gpurun_out
timeout 300 python tools/sweep.py --scenes dense_monkey --spp 16 --variants 4,15 --blocks 128 --reps 1 > gpurun_out/plain_tri2.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_render -c 4 -o gpurun_out/tri2_full -f python tools/sweep.py --scenes dense_monkey --spp 16 --variants 4,15 --blocks 128 --reps 1 > gpurun_out/ncu_tri2.log 2>&1
tail -3 gpurun_out/plain_tri2.log
