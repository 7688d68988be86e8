#!/bin/bash
# Shared frame buffers: the GPU test (attach + CUDA IPC), then bench.py --shard p2p next to --shard tiles on N GPUs.
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "shared_frame or tile_sharding" > gpurun_out/pytest_p2p.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_p2p.log
for mode in p2p tiles; do
  timeout 600 python bench.py --gpus $N --steps 3 --warmup 3 --shard $mode > gpurun_out/bench_n${N}_${mode}.json 2> gpurun_out/bench_n${N}_${mode}.err; echo rc=$?
  DRT_BENCH_SPP=1 timeout 600 python bench.py --gpus $N --steps 50 --warmup 10 --shard $mode > gpurun_out/bench_n${N}_${mode}_1spp.json 2> gpurun_out/bench_n${N}_${mode}_1spp.err; echo rc=$?
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/bench_n*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, d["n_gpus"], d["scaling"], "Mrays/s %.0f" % d["value"], "ms/step %.3f" % d["ms_per_step"], "e2e %.0f" % d["e2e"]["value"])
    except Exception as e:
        print(f, "unreadable", e)
PY
tail -5 gpurun_out/bench_n${N}_p2p.err
