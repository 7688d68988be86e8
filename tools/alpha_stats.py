#!/usr/bin/env python3
"""Per-ray traversal statistics of a scene (STATS build): node visits, triangle tests, alpha tests, ties.  python tools/alpha_stats.py [scene ...]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import dustraytracer_b200 as drt  # noqa: E402
import oracles as O  # noqa: E402

for name in sys.argv[1:] or ["animemaid", "valorant_ghost", "dense_monkey", "room"]:
    s = drt.Scene()
    s.loadGLTFmodel(O.scene_path(name))
    drt.BVHBuilder().buildIterative(s)
    r = drt.Renderer(flags=drt.DRT_FLAG_STATS)
    r.m_RendererSettings = drt.RendererSettings(enable_sunlight=1, ray_bounce_limit=2, max_samples=1 << 30)
    r.ResizeBuffer(1920, 1080)
    r.Render(drt.Camera(), s, 4)
    c = r.counters()
    rays = c["closest_rays"] + c["shadow_rays"]
    st = s.stats()
    print(json.dumps({"scene": name, "rays": rays, "nodes_per_ray": round(c["node_visits"] / rays, 3), "tris_per_ray": round(c["tri_tests"] / rays, 3),
                      "alpha_tests_per_ray": round(c["alpha_tests"] / rays, 4), "ties_per_Mray": round(1e6 * c["tie_events"] / rays, 2), "max_stack": c["max_stack"],
                      "alpha_tris": st["alpha_tris"], "alpha_opaque": st["alpha_opaque"], "alpha_hidden": st["alpha_hidden"]}), flush=True)
