#!/bin/bash
# Multi-GPU bench runs as the driver launches them. Usage: tools/gpu_scale.sh N [extra bench args]
N=$1; shift
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 20 --warmup 5 "$@" > gpurun_out/bench_n${N}.json 2> gpurun_out/bench_n${N}.err; echo "rc=$?" >> gpurun_out/bench_n${N}.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --impl reference --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_ref_n${N}.json 2> gpurun_out/bench_ref_n${N}.err; echo "rc=$?" >> gpurun_out/bench_ref_n${N}.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29535 bench.py --gpus $N --steps 20 --warmup 5 --scaling weak > gpurun_out/bench_n${N}_weak.json 2> gpurun_out/bench_n${N}_weak.err; echo "rc=$?" >> gpurun_out/bench_n${N}_weak.err
tail -n 3 gpurun_out/bench_n${N}.err; tail -n 3 gpurun_out/bench_ref_n${N}.err
cat gpurun_out/bench_n${N}.json gpurun_out/bench_ref_n${N}.json gpurun_out/bench_n${N}_weak.json | cut -c1-400
