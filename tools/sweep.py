#!/usr/bin/env python3
"""Quick device-side timing of the render kernel over scenes / kernel variants / block sizes (one GPU).
    python tools/sweep.py [--scenes room,dense_monkey] [--spp 32] [--variants 0,1] [--blocks 64,128,256] [--reps 3]
Prints one JSON line per combination (Mrays/s from the kernel's own ray counters and CUDA-event time)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import dustraytracer_b200 as drt  # noqa: E402

CONFIGS = {  # BASELINE.json configs: (sun, bounces)
    "room": (1, 8), "dense_monkey": (0, 4), "valorant_ghost": (1, 2), "animemaid": (1, 2), "cornell_box": (0, 2),
    "mc_transparency_test": (1, 2),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scenes", default="room")
    ap.add_argument("--spp", type=int, default=32)
    ap.add_argument("--variants", default="0,1", help="kernel variants (drt.h); -1 = automatic selection (launch auto-tuner)")
    ap.add_argument("--blocks", default="64,128,256")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--flags", type=int, default=0)
    ap.add_argument("--calls", type=int, default=1, help="render calls per measurement, --spp samples each (1 = one batched call)")
    ap.add_argument("--bias", default="0:0:0", help="comma list of node:tri:shade vote biases (phased variants)")
    a = ap.parse_args()
    for name in a.scenes.split(","):
        sun, bounces = CONFIGS.get(name, (1, 2))
        scene = drt.Scene()
        scene.loadGLTFmodel(os.path.join(ROOT, "tests", "golden", "scenes", name + ".glb"))
        drt.BVHBuilder().buildIterative(scene)
        st = scene.stats()
        cam = drt.Camera()
        for v in [int(x) for x in a.variants.split(",")]:
            for b, bias in [(int(x), y) for x in a.blocks.split(",") for y in (a.bias.split(",") if v >= 2 else ["0:0:0"])]:
                os.environ["DRT_SCHED_BIAS"] = bias.replace(":", ",")
                r = drt.Renderer(flags=a.flags)
                r.m_RendererSettings = drt.RendererSettings(enable_sunlight=sun, ray_bounce_limit=bounces, max_samples=1 << 30)
                r.set_tuning(v, b)
                r.ResizeBuffer(a.width, a.height)
                r.Render(cam, scene, min(a.spp, 4))  # warm-up
                best = None
                for _ in range(a.reps):
                    r.resetAccumulationBuffer()
                    r.counters(reset=True)
                    ms = sum(r.Render(cam, scene, a.spp) for _ in range(a.calls))
                    c = r.counters()
                    rays = c["closest_rays"] + c["shadow_rays"]
                    if best is None or ms < best[0]:
                        best = (ms, rays, c)
                ms, rays, c = best
                out = {"scene": name, "tris": st["triangles"], "wide_nodes": st["wide_nodes"], "build_ms": round(st["build_ms"], 3),
                       "variant": v, "block": b, "bias": bias, "spp": a.spp, "calls": a.calls, "ms": round(ms, 3), "ms_per_sample": round(ms / (a.spp * a.calls), 4), "Mrays_s": round(rays / ms / 1e3, 1),
                       "Msamples_s": round(a.width * a.height * a.spp * a.calls / ms / 1e3, 1), "rays": rays}
                if a.flags & 4:
                    out.update({"nodes_per_ray": round(c["node_visits"] / rays, 3), "tris_per_ray": round(c["tri_tests"] / rays, 3),
                                "max_stack": c["max_stack"]})
                print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
