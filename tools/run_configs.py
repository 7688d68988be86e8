#!/usr/bin/env python3
"""Runs the five BASELINE.json configs on ONE GPU through the C ABI and prints one JSON line per config:
build time, Mrays/s, samples/s, wide-node visits / triangle tests per ray (STATS build, 1 spp), and -- for the
bundled scenes -- the reference's own CUDA path on the same GPU (oracle/_ref/libdrt_refcuda_fast.so, when present).

    python tools/run_configs.py [--configs 1,2,3,4a,4b,5s,5m,5] [--spp-scale 1.0] [--ref-frames 4]
"""
import argparse
import ctypes
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import dustraytracer_b200 as drt  # noqa: E402

SCENES = os.path.join(ROOT, "tests", "golden", "scenes")
# name -> (scene, W, H, spp, bounces, sun)   
CONFIGS = {
    "1": ("cornell_box", 512, 512, 1, 0, 0),
    "2": ("dense_monkey", 1920, 1080, 64, 4, 0),
    "3": ("room", 1920, 1080, 256, 8, 1),
    "4a": ("valorant_ghost", 1920, 1080, 128, 2, 1),
    "4b": ("animemaid", 1920, 1080, 128, 2, 1),
    "5s": ("soup:100000", 3840, 2160, 16, 4, 1),
    "5m": ("soup:1000000", 3840, 2160, 16, 4, 1),
    "5": ("soup:10000000", 3840, 2160, 64, 4, 1),
}


def pcg_hash(x):
    x = x.astype(np.uint32)
    with np.errstate(over="ignore"):
        s = x * np.uint32(747796405) + np.uint32(2891336453)
        w = ((s >> ((s >> np.uint32(28)) + np.uint32(4))) ^ s) * np.uint32(277803737)
    return (w >> np.uint32(22)) ^ w


def soup(n, seed=0x0D057A11):
    """counter-based generator on the reference's pcg_hash so C++/CUDA/Python agree bit for bit."""
    out = np.empty((n, 9), np.float32)
    chunk = 1 << 20
    for a in range(0, n, chunk):
        i = np.arange(a, min(n, a + chunk), dtype=np.uint32)
        base = pcg_hash(np.uint32(seed) ^ i)
        with np.errstate(over="ignore"):
            r = np.stack([pcg_hash(base * np.uint32(12) + np.uint32(k)) for k in range(12)], 1).astype(np.float64) / 4294967296.0
        c = r[:, 0:3] * 100.0 - 50.0
        for j in range(3):
            out[a:a + len(i), 3 * j:3 * j + 3] = (c + (r[:, 3 + 3 * j:6 + 3 * j] - 0.5) * 0.5).astype(np.float32)
    return out


def load_scene(name):
    t0 = time.perf_counter()
    if name.startswith("soup:"):
        scene = drt.Scene.from_triangles(soup(int(name.split(":")[1])))  # one material, albedo (0.8, 0.8, 0.8)
    else:
        scene = drt.Scene()
        scene.loadGLTFmodel(os.path.join(SCENES, name + ".glb"))
    load_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    drt.BVHBuilder().buildIterative(scene)
    return scene, load_s, time.perf_counter() - t0


def ref_cuda(name, cam_bytes, st_bytes, W, H, frames):
    path = os.path.join(ROOT, "oracle", "_ref", "libdrt_refcuda_fast.so")
    if name.startswith("soup:") or not os.path.exists(path):
        return None
    L = ctypes.CDLL(path)
    L.refcuda_load.restype = ctypes.c_void_p
    L.refcuda_render.restype = ctypes.c_double
    saved = os.dup(1)
    os.dup2(2, 1)
    try:
        h = L.refcuda_load(os.path.join(SCENES, name + ".glb").encode(), 8, 20)
    finally:
        os.dup2(saved, 1)
        os.close(saved)
    P = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    L.refcuda_render(ctypes.c_void_p(h), P(cam_bytes), P(st_bytes), W, H, 1, 1, None)
    ms = L.refcuda_render(ctypes.c_void_p(h), P(cam_bytes), P(st_bytes), W, H, 1, frames, None)
    L.refcuda_free(ctypes.c_void_p(h))
    return ms / frames


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="1,2,3,4a,4b,5s,5m")
    ap.add_argument("--spp-scale", type=float, default=1.0)
    ap.add_argument("--ref-frames", type=int, default=4)
    ap.add_argument("--variant", type=int, default=-1)
    ap.add_argument("--block", type=int, default=0, help="threads per block (0 = default)")
    a = ap.parse_args()
    import oracles as O
    for key in a.configs.split(","):
        name, W, H, spp, bounces, sun = CONFIGS[key]
        spp = max(1, int(round(spp * a.spp_scale)))
        scene, load_s, build_wall_s = load_scene(name)
        st = scene.stats()
        cam = drt.Camera(pos=(0.0, 0.0, 120.0)) if name.startswith("soup:") else drt.Camera()
        cam_bytes = O.default_camera_bytes()
        if name.startswith("soup:"):
            cam_bytes[8:11] = (0, 0, 120)
        out = {"config": key, "scene": name, "width": W, "height": H, "spp": spp, "bounces": bounces, "sun": sun,
               "tris": st["triangles"], "wide_nodes": st["wide_nodes"], "bvh_levels": st["bvh_levels"], "build_ms": round(st["build_ms"], 3),
               "build_wall_s": round(build_wall_s, 3), "load_s": round(load_s, 3), "device_MB": round(st["device_bytes"] / 1e6, 2)}
        settings = drt.RendererSettings(enable_sunlight=sun, ray_bounce_limit=bounces, max_samples=1 << 30)
        # traversal statistics (1 spp, STATS build)
        rs = drt.Renderer(flags=drt.DRT_FLAG_STATS)
        rs.m_RendererSettings = settings
        rs.ResizeBuffer(W, H)
        rs.Render(cam, scene, 1)
        c = rs.counters()
        rays1 = c["closest_rays"] + c["shadow_rays"]
        out.update({"nodes_per_ray": round(c["node_visits"] / rays1, 3), "tris_per_ray": round(c["tri_tests"] / rays1, 3),
                    "max_stack": c["max_stack"], "rays_per_sample": rays1})
        del rs
        r = drt.Renderer()
        r.m_RendererSettings = settings
        if a.variant >= 0 or a.block:
            r.set_tuning(a.variant if a.variant >= 0 else 4, a.block)
        r.ResizeBuffer(W, H)
        for _ in range(8 if spp <= 8 else 1):  # warm-up (few-sample launches: also lets the launch auto-tuner settle)
            r.Render(cam, scene, min(spp, 2))
        r.resetAccumulationBuffer()
        r.counters(reset=True)
        ms = r.Render(cam, scene, spp)
        c = r.counters()
        rays = c["closest_rays"] + c["shadow_rays"]
        img = r.readback(drt.DRT_RGBA32F)
        ordered_first = r.last_launch_ordered()
        r.resetAccumulationBuffer()
        ms_repeat = r.Render(cam, scene, spp)   # the same view again: blocks in the cost order measured by the launch before
        out.update({"ms_repeat": round(ms_repeat, 3), "Mrays_s_repeat": round(rays / ms_repeat / 1e3, 1), "ordered": [ordered_first, r.last_launch_ordered()],
                    "variant": r.last_variant()})
        out.update({"ms": round(ms, 3), "Mrays_s": round(rays / ms / 1e3, 1), "Msamples_s": round(W * H * spp / ms / 1e3, 1),
                    "rays": rays, "mean_ldr": round(float(img[:, :, :3].mean()), 5), "finite": bool(np.isfinite(img).all())})
        ref_ms = ref_cuda(name, cam_bytes, O.settings_bytes(sun=bool(sun), bounces=bounces, max_samples=1 << 30), W, H, a.ref_frames) if a.ref_frames > 0 else None
        if ref_ms:
            out["ref_cuda_fast_ms_per_sample"] = round(ref_ms, 4)
            out["speedup_vs_ref_cuda_fast"] = round(ref_ms / (ms / spp), 2)
        print(json.dumps(out), flush=True)
        del r, scene


if __name__ == "__main__":
    main()
