#!/bin/bash
# Round-2 GPU session pieces. Usage: tools/gpu_r02.sh [tests] [bench] [ref] [launches] [ncu256] [ncufull16] ...
mkdir -p gpurun_out
METRICS=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum,l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum,l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum,dram__bytes_read.sum,dram__bytes_write.sum,l1tex__data_pipe_lsu_wavefronts.sum,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active
for what in "$@"; do
case $what in
tests)
  timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
  timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log ;;
ref)
  timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?" >> gpurun_out/bench_ref.err ;;
bench)
  timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err ;;
launches)
  DRT_BENCH_SKIP_BASELINES=1 timeout 900 python bench.py --steps 2 --warmup 3 > gpurun_out/plain.log 2>&1 &&
  DRT_BENCH_SKIP_BASELINES=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_launch.log 2>&1 ;;
ncu256)
  # the benchmarked launch itself (256 spp), short metric list
  DRT_BENCH_SKIP_BASELINES=1 timeout 600 python bench.py --steps 1 --warmup 3 > gpurun_out/plain256.log 2>&1 &&
  DRT_BENCH_SKIP_BASELINES=1 timeout 1200 ncu --metrics $METRICS --clock-control none -k regex:k_render --launch-skip 3 -c 1 -o gpurun_out/render256 -f python bench.py --steps 1 --warmup 3 > gpurun_out/ncu256.log 2>&1 ;;
ncufull16)
  DRT_BENCH_SKIP_BASELINES=1 DRT_BENCH_SPP=16 timeout 600 python bench.py --steps 1 --warmup 3 > gpurun_out/plain16.log 2>&1 &&
  DRT_BENCH_SKIP_BASELINES=1 DRT_BENCH_SPP=16 timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_render -c 1 -o gpurun_out/render_full -f python bench.py --steps 1 --warmup 3 > gpurun_out/ncu_full.log 2>&1 ;;
ncurg)
  # the regrouping kernel (variant 16), full set on a 16-spp launch
  DRT_VARIANT=16 DRT_BENCH_SKIP_BASELINES=1 DRT_BENCH_SPP=16 timeout 600 python bench.py --steps 1 --warmup 3 > gpurun_out/plain_rg.log 2>&1 &&
  DRT_VARIANT=16 DRT_BENCH_SKIP_BASELINES=1 DRT_BENCH_SPP=16 timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_render_regroup -c 1 -o gpurun_out/render_rg -f python bench.py --steps 1 --warmup 3 > gpurun_out/ncu_rg.log 2>&1 ;;
sweep)
  timeout 900 python tools/sweep.py $SWEEP_ARGS > gpurun_out/sweep.jsonl 2> gpurun_out/sweep.err ;;
*)
  echo "unknown piece $what" ;;
esac
done
ls -la gpurun_out | tail -20
