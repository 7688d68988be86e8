#!/bin/bash
# Pixel-queue kernel and launch auto-tuner: parity first, then one-sample-launch timings: variant 4, variant 9, automatic (-1).
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "pixel_queue or bit_exact or full_size or autotuner" > gpurun_out/pytest_queue.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_queue.log
tail -5 gpurun_out/pytest_queue.log
: > gpurun_out/sweep_queue.jsonl
S=room,dense_monkey,valorant_ghost,animemaid,cornell_box
for spp in 1 2 4 8; do
  calls=$((256 / spp))
  timeout 600 python tools/sweep.py --scenes $S --spp $spp --calls $calls --variants 4,-1 --blocks 128 --reps 2 >> gpurun_out/sweep_queue.jsonl 2>> gpurun_out/sweep_queue.err
done
timeout 600 python tools/sweep.py --scenes $S --spp 64 --calls 1 --variants 4,-1 --blocks 128 --reps 2 >> gpurun_out/sweep_queue.jsonl 2>> gpurun_out/sweep_queue.err
python - <<'PY'
import json, collections
t = collections.OrderedDict()
for l in open("gpurun_out/sweep_queue.jsonl"):
    d = json.loads(l)
    t.setdefault((d["scene"], d["spp"]), []).append(("v%d" % d["variant"], d["ms_per_sample"]))
for k, v in t.items():
    print(k[0], "spp", k[1], " ".join("%s=%.4f" % x for x in v), "auto/v4 speed-up %.3f" % (v[0][1] / v[1][1]))
PY
