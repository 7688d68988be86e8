#!/usr/bin/env python3
"""The reference's call pattern: Renderer::Render once per displayed frame = ONE sample per launch (Renderer.cu:80-117).
Times N such calls through the C ABI (wall clock, each call returns with the frame complete) -- with the renderer's
automatic kernel selection (launch auto-tuner: fixed-pixel kernel or persistent pixel-queue kernel, whichever it measures
faster) and with the fixed-pixel kernel forced -- next to one batched call and to the reference's own CUDA path launched
the same way."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import dustraytracer_b200 as drt  # noqa: E402
import bench  # noqa: E402

N = 512
scene = drt.Scene()
scene.loadGLTFmodel(bench.SCENE_PATH)
drt.BVHBuilder().buildIterative(scene)
cam = drt.Camera()


def one_sample_calls(variant):
    r = drt.Renderer()
    r.set_tuning(variant, 0)
    r.m_RendererSettings = drt.RendererSettings(enable_sunlight=1, ray_bounce_limit=8, max_samples=1 << 30)
    r.ResizeBuffer(1920, 1080)
    for _ in range(8):
        r.Render(cam, scene)
    r.resetAccumulationBuffer()
    r.counters(reset=True)
    t0 = time.perf_counter()
    dev_ms = 0.0
    for _ in range(N):   # (includes the auto-tuner's probe launches: 6 of every 256)
        dev_ms += r.Render(cam, scene)
    wall = time.perf_counter() - t0
    c = r.counters(reset=True)
    rays = c["closest_rays"] + c["shadow_rays"]
    return r, {"wall_ms_per_sample": wall / N * 1e3, "device_ms_per_sample": dev_ms / N, "Mrays_s_wall": rays / wall / 1e6}


r, auto = one_sample_calls(-1)
_, fixed = one_sample_calls(4)
r.resetAccumulationBuffer()
t0 = time.perf_counter()
ms_b = r.Render(cam, scene, N)
wall_b = time.perf_counter() - t0
bench.claim_stdout()
ref = bench.ref_cuda_baseline(frames=64)
bench.emit_json_line({"scene": "room.glb 1920x1080, 8 bounces, sun", "calls": N,
                      "one_sample_per_call": auto, "one_sample_per_call_fixed_pixel_kernel": fixed,
                      "one_batched_call": {"wall_ms_per_sample": wall_b / N * 1e3, "device_ms_per_sample": ms_b / N},
                      "reference_cuda_one_sample_per_launch_ms": {k: v["ms_per_sample"] for k, v in (ref or {}).items()}})
