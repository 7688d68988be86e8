#!/bin/bash
# A/B of measurement builds (dustraytracer_b200/ab/libdrt_<name>.so) against the in-tree library. Usage: tools/gpu_ab_libs.sh name1 name2 ...
mkdir -p gpurun_out
: > gpurun_out/ab_libs.jsonl
for name in base "$@"; do
  if [ $name = base ]; then unset DRT_LIB_PATH; else export DRT_LIB_PATH=$PWD/dustraytracer_b200/ab/libdrt_$name.so; fi
  echo "# lib $name" >> gpurun_out/ab_libs.jsonl
  timeout 600 python tools/sweep.py --scenes room --spp 256 --variants -1 --blocks 128 --reps 3 2>/dev/null | grep "^{" >> gpurun_out/ab_libs.jsonl
  timeout 600 python tools/sweep.py --scenes cornell_box,dense_monkey,valorant_ghost,animemaid --spp 64 --variants -1 --blocks 128 --reps 3 2>/dev/null | grep "^{" >> gpurun_out/ab_libs.jsonl
  timeout 600 python tools/run_configs.py --configs 5s --ref-frames 0 2>/dev/null | grep "^{" >> gpurun_out/ab_libs.jsonl
done
python - <<'PY'
import json
for l in open('gpurun_out/ab_libs.jsonl'):
    l=l.strip()
    if l.startswith('#'): print(l); continue
    d=json.loads(l)
    print("  %-16s nodes %7d  %9.3f ms  %9.1f Mrays/s  V/T %s %s" % (d['scene'], d['wide_nodes'], d['ms'], d['Mrays_s'], d.get('nodes_per_ray'), d.get('tris_per_ray')))
PY
