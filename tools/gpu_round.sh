#!/bin/bash
# One GPU session. Usage: tools/gpu_round.sh [tests] [sweep] [bench] [ncu] [ncuq] [dropin]
set -x
mkdir -p gpurun_out
for what in "$@"; do
case $what in
tests)
  timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
  timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log ;;
sweep)
  timeout 900 python tools/sweep.py $SWEEP_ARGS > gpurun_out/sweep.jsonl 2> gpurun_out/sweep.err ;;
bench)
  timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
  timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err ;;
dropin)
  timeout 600 python tools/dropin_mode.py > gpurun_out/dropin_mode.json 2> gpurun_out/dropin_mode.err ;;
ncu)
  timeout 900 python bench.py --steps 2 --warmup 3 > gpurun_out/plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_launch.log 2>&1
  DRT_BENCH_SPP=16 timeout 600 python bench.py --steps 1 --warmup 3 > gpurun_out/plain2.log 2>&1 &&
  DRT_BENCH_SPP=16 timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_render -c 1 -o gpurun_out/render_full -f python bench.py --steps 1 --warmup 3 > gpurun_out/ncu_full.log 2>&1 ;;
ncuq)
  # the persistent pixel-queue kernel on a one-sample launch (the reference's interactive call pattern)
  DRT_BENCH_SPP=1 DRT_VARIANT=9 timeout 600 python bench.py --steps 1 --warmup 3 > gpurun_out/plain3.log 2>&1 &&
  DRT_BENCH_SPP=1 DRT_VARIANT=9 timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_render_cold -c 1 -o gpurun_out/render_queue_full -f python bench.py --steps 1 --warmup 3 > gpurun_out/ncu_queue_full.log 2>&1 ;;
esac
done
