#!/bin/bash
# Sample-synchronised BEGIN step: parity, then batched launches (64 spp) over the begin threshold, then few-sample launches.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "pixel_queue or bit_exact or full_size or autotuner" > gpurun_out/pytest_queue.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_queue.log
tail -5 gpurun_out/pytest_queue.log
: > gpurun_out/sweep_queue.jsonl
S=room,dense_monkey,valorant_ghost,animemaid,cornell_box
B=0:1:0,0:8:0,0:12:0,0:16:0,0:20:0,0:24:0
timeout 600 python tools/sweep.py --scenes $S --spp 64 --calls 1 --variants 4 --blocks 128 --reps 2 >> gpurun_out/sweep_queue.jsonl 2>> gpurun_out/sweep_queue.err
timeout 900 python tools/sweep.py --scenes $S --spp 64 --calls 1 --variants 12,11,9 --blocks 128 --reps 2 --bias $B >> gpurun_out/sweep_queue.jsonl 2>> gpurun_out/sweep_queue.err
for spp in 1 4; do
  calls=$((64 / spp))
  timeout 600 python tools/sweep.py --scenes $S --spp $spp --calls $calls --variants 4 --blocks 128 --reps 2 >> gpurun_out/sweep_queue.jsonl 2>> gpurun_out/sweep_queue.err
  timeout 600 python tools/sweep.py --scenes $S --spp $spp --calls $calls --variants 9 --blocks 128 --reps 2 --bias 0:4:0,0:10:0,0:16:0 >> gpurun_out/sweep_queue.jsonl 2>> gpurun_out/sweep_queue.err
done
python - <<'PY'
import json, collections
t = collections.OrderedDict()
for l in open("gpurun_out/sweep_queue.jsonl"):
    d = json.loads(l)
    t.setdefault((d["scene"], d["spp"]), []).append(("v%d/%s" % (d["variant"], d["bias"].split(":")[1]), d["ms_per_sample"]))
for k, v in t.items():
    print(k[0], "spp", k[1], " ".join("%s=%.4f" % x for x in v))
PY
