#!/usr/bin/env python3
"""Design tool: warp-schedule simulator for the shipped three-step kernel (NODE / TRI / SHADE, full vote + sticky
re-issue; render_warp_phased in drt_phased.cuh). CPU only.

tests/hostemu emu_warp_sim3 runs one warp (an 8x4 pixel tile) with the REAL step functions under a vote policy and
counts issues and lane-steps per step. This script sweeps policies and prints modelled warp instructions per ray:
    cost = NODE issues * cN + TRI issues * cT + SHADE issues * cS + (worst rejection-loop draws per SHADE issue) * cD
           + full votes * cV + sticky ballots * cB + (SHADE issues in which a lane ends its sample + END issues) * cE

    python tools/sched_sim3.py [--scene room|dense_monkey|...|soup:100000] [--tiles 40] [--spp 8]
"""
import argparse
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import oracles as O  # noqa: E402

# name, weights N T S, eager thresholds N T S, sticky N T S
POLICIES = [
    ("shipped: argmax, stick 6", (1, 1, 1), (0, 0, 0), (6, 6, 6)),
    ("argmax, no stick", (1, 1, 1), (0, 0, 0), (33, 33, 33)),
    ("stick 4", (1, 1, 1), (0, 0, 0), (4, 4, 4)),
    ("stick 10", (1, 1, 1), (0, 0, 0), (10, 10, 10)),
    ("stick N6 T6 S12", (1, 1, 1), (0, 0, 0), (6, 6, 12)),
    ("stick N6 T3 S12", (1, 1, 1), (0, 0, 0), (6, 3, 12)),
    ("stick N8 T2 S33", (1, 1, 1), (0, 0, 0), (8, 2, 33)),
    ("TRI eager>=1", (1, 1, 1), (0, 1, 0), (6, 1, 6)),
    ("TRI eager>=4", (1, 1, 1), (0, 4, 0), (6, 4, 6)),
    ("TRI eager>=8", (1, 1, 1), (0, 8, 0), (6, 6, 6)),
    ("TRI>=4, SHADE>=4 eager", (1, 1, 1), (0, 4, 4), (6, 4, 4)),
    ("TRI>=4, SHADE>=8 eager", (1, 1, 1), (0, 4, 8), (6, 4, 6)),
    ("TRI>=2, SHADE>=2 eager", (1, 1, 1), (0, 2, 2), (6, 2, 2)),
    ("w T2 S2", (1, 2, 2), (0, 0, 0), (6, 6, 6)),
    ("w T2 S1", (1, 2, 1), (0, 0, 0), (6, 6, 6)),
    ("w T3 S3", (1, 3, 3), (0, 0, 0), (6, 4, 4)),
    ("w T1 S0.5 (shade late)", (1, 1, 0.5), (0, 0, 0), (6, 6, 12)),
    ("w T2 S0.5", (1, 2, 0.5), (0, 0, 0), (6, 4, 12)),
    ("w T4 S0.5", (1, 4, 0.5), (0, 0, 0), (6, 2, 12)),
    ("w N0.5 T1 S0.5", (0.5, 1, 0.5), (0, 0, 0), (12, 4, 12)),
    # no vote: every iteration runs NODE, TRI, SHADE in turn for the lanes that want them (thresholds N T S)
    ("cycle 1/1/1", (1, 1, 1), (1, 1, 1), (33, 33, 33), 1),
    ("cycle 1/1/4", (1, 1, 1), (1, 1, 4), (33, 33, 33), 1),
    ("cycle 1/1/8", (1, 1, 1), (1, 1, 8), (33, 33, 33), 1),
    ("cycle 1/1/12", (1, 1, 1), (1, 1, 12), (33, 33, 33), 1),
    ("cycle 4/4/8", (1, 1, 1), (4, 4, 8), (33, 33, 33), 1),
    ("cycle 6/6/12", (1, 1, 1), (6, 6, 12), (33, 33, 33), 1),
    ("cycle 8/4/12", (1, 1, 1), (8, 4, 12), (33, 33, 33), 1),
    # four steps: the sample end (post-process, accumulate, next camera ray) leaves SHADE and is voted on its own (weight)
    ("4-step END w1", (1, 1, 1), (0, 0, 0), (6, 6, 6), 0, 1.0),
    ("4-step END w0.5", (1, 1, 1), (0, 0, 0), (6, 6, 6), 0, 0.5),
    ("4-step END w0.25", (1, 1, 1), (0, 0, 0), (6, 6, 6), 0, 0.25),
    ("4-step END w0.7", (1, 1, 1), (0, 0, 0), (6, 6, 6), 0, 0.7),
]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default="room")
    ap.add_argument("--tiles", type=int, default=40)
    ap.add_argument("--spp", type=int, default=8)
    ap.add_argument("--sun", type=int, default=1)
    ap.add_argument("--bounces", type=int, default=8)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--cost", default="132,71,70,15,16,10,190", help="cN,cT,cS,cD,cV,cB,cE")
    a = ap.parse_args()
    cN, cT, cS, cD, cV, cB, cE = [float(x) for x in a.cost.split(",")]
    emu = O.HostEmu()
    emu.L.emu_warp_sim3.restype = None
    cam = O.default_camera_bytes()
    if a.scene.startswith("soup:"):
        from run_configs import soup
        mat = np.zeros((1, 11), np.uint32)
        mat.view(np.float32)[0, 0:3] = 0.8
        mat.view(np.int32)[0, 6] = -1
        h = emu.from_triangles(soup(int(a.scene.split(":")[1])), mat)
        cam[8:11] = (0, 0, 120)
    else:
        h = emu.load(O.scene_path(a.scene))
    st = O.settings_bytes(sun=bool(a.sun), bounces=a.bounces, max_samples=1 << 30)
    P = lambda x: x.ctypes.data_as(ctypes.c_void_p)
    W, H = a.width, a.height
    rng = np.random.default_rng(1)
    tiles = [(int(rng.integers(0, W // 8)), int(rng.integers(0, H // 4))) for _ in range(a.tiles)]
    print("%-40s %9s %7s  %-22s %-17s %s" % ("policy", "winst/ray", "vs ship", "lanes/issue N T S", "issues/ray N T S", "votes+ballots/ray"))
    base = None
    for entry in POLICIES:
        name, w, eager, stick = entry[:4]
        cycle = entry[4] if len(entry) > 4 else 0
        w_end = entry[5] if len(entry) > 5 else 0.0
        out = (ctypes.c_ulonglong * 13)()
        pol = np.array(list(w) + list(eager) + list(stick) + [cycle, 1.0 if w_end > 0 else 0.0, w_end], np.float32)
        for tx, ty in tiles:
            emu.L.emu_warp_sim3(h, P(cam), P(st), W, H, tx, ty, 1, a.spp, P(pol), out)
        ex = np.array(out[0:3], float)
        ln = np.array(out[3:6], float)
        votes, ballots, rays, worst = float(out[6]), float(out[7]), float(out[8]), float(out[9])
        ended_issues, end_issues, end_lanes = float(out[10]), float(out[11]), float(out[12])
        if cycle:  # one loop iteration = three "does any lane want it" branches + one count, ~12 instructions
            cost = ex[0] * cN + ex[1] * cT + ex[2] * cS + worst * cD + votes * 12.0
        else:      # a SHADE issue in which some lane ends its sample also runs the sample end (cE) for those lanes
            cost = ex[0] * cN + ex[1] * cT + ex[2] * cS + worst * cD + votes * cV + ballots * cB + (ended_issues + end_issues) * cE
        if end_issues:
            name += " (%.1f lanes/END)" % (end_lanes / end_issues)
        per_ray = cost / rays
        if base is None:
            base = per_ray
        print("%-40s %9.1f %7.3f  %-22s %-17s %.3f" % (name, per_ray, per_ray / base, np.round(ln / np.maximum(ex, 1), 1),
                                                    np.round(ex / rays, 3), (votes + ballots) / rays))


if __name__ == "__main__":
    main()
