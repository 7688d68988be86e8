#!/usr/bin/env python3
"""Cost-ordered launches against row-major launches (DRT_FLAG_NO_TILE_ORDER), same process, same box: device ms of the 2nd .. 6th
batched launch of a view (the first one records the costs), each from a reset accumulation buffer.  python tools/order_ab.py"""
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import dustraytracer_b200 as drt  # noqa: E402
import oracles as O  # noqa: E402

CASES = [tuple(int(x) if x.lstrip("-").isdigit() else x for x in c.split(":")) for c in os.environ["DRT_ORDER_CASES"].split(",")] if os.environ.get("DRT_ORDER_CASES") else [("room", 256, 8, 1), ("room", 32, 8, 1), ("cornell_box", 64, 8, 1), ("dense_monkey", 64, 4, 0), ("valorant_ghost", 128, 2, 1), ("animemaid", 128, 2, 1),
         ("sunshadow_test", 64, 8, 1), ("mc_transparency_test", 64, 8, 1)]
for name, spp, bounces, sun in CASES:
    s = drt.Scene()
    s.loadGLTFmodel(O.scene_path(name))
    drt.BVHBuilder().buildIterative(s)
    row = {"scene": name, "spp": spp, "width": 1920, "height": 1080}
    for label, flags in (("row_major", drt.DRT_FLAG_NO_TILE_ORDER), ("cost_order", 0)):
        r = drt.Renderer(flags=flags)
        r.m_RendererSettings = drt.RendererSettings(enable_sunlight=sun, ray_bounce_limit=bounces, max_samples=1 << 30)
        r.ResizeBuffer(1920, 1080)
        cam = drt.Camera()
        ms = []
        first = None
        for k in range(7):
            r.resetAccumulationBuffer()
            t = r.Render(cam, s, spp)
            if k == 0:
                first = t   # (cost order: one-sample probe launch + sort + the rest in cost order when spp >= 64)
            if k >= 2:
                ms.append(t)
        row[label + "_first_launch_ms"] = round(first, 3)
        c = r.counters(reset=True)
        row[label + "_ms"] = round(statistics.median(ms), 3)
        row[label + "_ordered"] = r.last_launch_ordered()
        row[label + "_variant"] = r.last_variant()
    row["gain_pct"] = round(100.0 * (row["row_major_ms"] / row["cost_order_ms"] - 1.0), 2)
    print(json.dumps(row), flush=True)
