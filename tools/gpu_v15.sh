#!/bin/bash
# Two-triangle TRI step (variant 15) against the default kernel (variant 4): parity, then timings.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "bit_exact or pixel_queue or full_size" > gpurun_out/pytest_v15.log 2>&1; tail -3 gpurun_out/pytest_v15.log
timeout 600 python tools/sweep.py --scenes room,dense_monkey,valorant_ghost,animemaid,cornell_box --spp 64 --variants ${V:-4,15} --blocks 128 --reps 3 > gpurun_out/sweep_v15.jsonl 2>/dev/null
timeout 300 python tools/sweep.py --scenes room --spp 256 --variants ${V:-4,15} --blocks 128 --reps 2 >> gpurun_out/sweep_v15.jsonl 2>/dev/null
python - <<'PY'
import json
for l in open("gpurun_out/sweep_v15.jsonl"):
    d = json.loads(l)
    print(d["scene"], "v%d" % d["variant"], "spp", d["spp"], d["ms_per_sample"], d["Mrays_s"])
PY
