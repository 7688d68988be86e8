#!/usr/bin/env python3
"""BASELINE config 3 through the in-library multi-GPU path (drt_multi_*: ONE process, no torch, NCCL loaded by libdrt.so):
one 256-spp 1080p frame of room.glb per step on 1, 2, 4, 8 devices, sample-range sharding (and image tiles with --tiles).
Time = wall clock around drt_multi_render (launches on all devices + grouped ncclReduce + resolve + stream syncs).

    python tools/multi_bench.py [--steps 10] [--spp 256] [--tiles]
"""
import argparse
import ctypes
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import dustraytracer_b200 as drt  # noqa: E402
from dustraytracer_b200 import _lib  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--spp", type=int, default=256)
    ap.add_argument("--tiles", action="store_true")
    a = ap.parse_args()
    n_avail = 1
    try:
        import subprocess
        n_avail = len(subprocess.run(["nvidia-smi", "-L"], stdout=subprocess.PIPE, text=True).stdout.strip().splitlines())
    except Exception:
        pass
    scene = drt.Scene()
    scene.loadGLTFmodel(os.path.join(ROOT, "tests", "golden", "scenes", "room.glb"))
    cam = drt.Camera()
    base = None
    for n in (1, 2, 4, 8):
        if n > n_avail:
            break
        mr = drt.MultiRenderer(list(range(n)))
        mr.set_scene(scene)
        mr.set_sharding(tiles=a.tiles)
        mr.m_RendererSettings = drt.RendererSettings(enable_sunlight=1, ray_bounce_limit=8, max_samples=1 << 30)
        mr.ResizeBuffer(1920, 1080)
        for _ in range(3):
            mr.resetAccumulationBuffer()
            mr.Render(cam, a.spp)
        mr.counters(reset=True)
        ms = []
        for _ in range(a.steps):
            mr.resetAccumulationBuffer()
            ms.append(mr.Render(cam, a.spp))
        c = mr.counters(reset=True)
        rays = (c["closest_rays"] + c["shadow_rays"]) / a.steps
        med = statistics.median(ms)
        base = base or med
        print(json.dumps({"devices": n, "sharding": "tiles" if a.tiles else "sample ranges", "spp_per_frame": a.spp, "ms_per_frame_median": round(med, 3),
                          "ms_per_frame_min": round(min(ms), 3), "Mrays_s": round(rays / med / 1e3, 1), "speedup_vs_1": round(base / med, 3),
                          "nccl_reduces": mr.nccl_reduces(), "what": "wall clock of drt_multi_render (one process, in-library NCCL)"}), flush=True)
        del mr


if __name__ == "__main__":
    main()
