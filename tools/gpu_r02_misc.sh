#!/bin/bash
# Round-2 side measurements: block size of the batched kernel, cost of the watertight TRI step, the BASELINE config table.
mkdir -p gpurun_out
{
echo "# block size, fixed-tile kernel (variant 4 room / 15 others), 1080p"
timeout 600 python tools/sweep.py --scenes room --spp 256 --variants 4 --blocks 32,64,128 --reps 3
timeout 600 python tools/sweep.py --scenes room --spp 32 --variants 4 --blocks 32,64,128 --reps 3
timeout 600 python tools/sweep.py --scenes dense_monkey,valorant_ghost,animemaid --spp 64 --variants 15 --blocks 32,64,128 --reps 3
echo "# watertight TRI step (flags 16 -> variant 21) against variant 4, 64 spp"
timeout 600 python tools/sweep.py --scenes room,dense_monkey,valorant_ghost,animemaid --spp 64 --variants 4 --blocks 128 --reps 3
timeout 600 python tools/sweep.py --scenes room,dense_monkey,valorant_ghost,animemaid --spp 64 --variants 4 --blocks 128 --reps 3 --flags 16
} > gpurun_out/misc_sweep.jsonl 2> gpurun_out/misc_sweep.err
timeout 1200 python tools/run_configs.py --configs 1,2,3,4a,4b,5s,5m,5 > gpurun_out/configs.jsonl 2> gpurun_out/configs.err; echo "configs rc=$?"
tail -n 5 gpurun_out/misc_sweep.err; tail -n 5 gpurun_out/configs.err
cut -c1-250 gpurun_out/misc_sweep.jsonl | tail -n 40
