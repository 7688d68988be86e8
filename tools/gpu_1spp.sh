#!/bin/bash
# One-sample launches on all scenes: fixed-pixel kernel forced (variant 4) against automatic selection (-1).
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -2 gpurun_out/pytest_gpu.log
timeout 600 python tools/sweep.py --scenes room,dense_monkey,valorant_ghost,animemaid,cornell_box --spp 1 --calls 256 --variants 4,-1 --blocks 128 --reps 2 > gpurun_out/sweep_1spp.jsonl 2>/dev/null
python - <<'PY'
import json
for l in open("gpurun_out/sweep_1spp.jsonl"):
    d = json.loads(l)
    print(d["scene"], "v%d" % d["variant"], "spp", d["spp"], d["ms_per_sample"], d["Mrays_s"])
PY
