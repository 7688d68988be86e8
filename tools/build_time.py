#!/usr/bin/env python
"""Warm BVH build times of the small bundled scenes: single-launch builder against the level-synchronous one.
Usage (GPU box): python tools/build_time.py [reps]"""
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import dustraytracer_b200 as drt  # noqa: E402
import oracles as O  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
for name in ("cube", "cornell_box", "room", "sunshadow_test", "mc_transparency_test"):
    s = drt.Scene()
    s.loadGLTFmodel(O.scene_path(name))
    row = {"scene": name, "block": os.environ.get("DRT_BUILD_SMALL_BLOCK", "auto")}
    for label, single in (("single_launch", True), ("multi_launch", False)):
        s.build_bvh(single_launch=single)  # warm-up (module load, allocator)
        ms = []
        for _ in range(reps):
            s.build_bvh(single_launch=single)
            ms.append(s.stats()["build_ms"])
        st = s.stats()
        row.update({"tris": st["triangles"], "bvh_levels": st["bvh_levels"], "wide_nodes": st["wide_nodes"],
                    label + "_ms_min": round(min(ms), 4), label + "_ms_median": round(sorted(ms)[len(ms) // 2], 4), label + "_launches": st["build_launches"]})
    print(json.dumps(row), flush=True)
