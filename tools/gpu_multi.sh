#!/bin/bash
# Both sharding modes of bench.py on N GPUs (N = $1): sample-range (weak scaling), image tiles (strong scaling), tiles at one sample.
N=${1:-8}
mkdir -p gpurun_out
timeout 600 python bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_n${N}_samples.json 2> gpurun_out/bench_n${N}_samples.err; echo rc=$?
timeout 600 python bench.py --gpus $N --steps 3 --warmup 3 --shard tiles > gpurun_out/bench_n${N}_tiles.json 2> gpurun_out/bench_n${N}_tiles.err; echo rc=$?
DRT_BENCH_SPP=1 timeout 600 python bench.py --gpus $N --steps 50 --warmup 10 --shard tiles > gpurun_out/bench_n${N}_tiles_1spp.json 2> gpurun_out/bench_n${N}_tiles_1spp.err; echo rc=$?
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/bench_n*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, d["n_gpus"], d["scaling"], "Mrays/s %.0f" % d["value"], "ms/step %.3f" % d["ms_per_step"], "e2e %.0f" % d["e2e"]["value"])
    except Exception as e:
        print(f, "unreadable", e)
PY
