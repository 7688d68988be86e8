#!/usr/bin/env python3
"""Design tool: warp-schedule simulator for the five-step render kernel (CPU only, no GPU needed).

tests/hostemu emu_warp_sim emulates one warp (an 8x4 pixel tile) with the REAL step functions of
drt_phased.cuh under a vote policy, and counts how often each step is issued and for how many lanes.
This script sweeps policies and prints issued warp-instructions per ray for a step-cost table.

    python tools/sched_sim.py [--scene room] [--tiles 40] [--spp 16] [--cost 200,90,60,95,190] [--vote 14]
"""
import argparse
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracles as O  # noqa: E402

STEPS = ["NODE", "TRI", "HIT", "UVEC", "END"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default="room")
    ap.add_argument("--tiles", type=int, default=40)
    ap.add_argument("--spp", type=int, default=16)
    ap.add_argument("--sun", type=int, default=1)
    ap.add_argument("--bounces", type=int, default=8)
    ap.add_argument("--cost", default="200,90,60,95,190", help="warp instructions per issued step")
    ap.add_argument("--vote", type=float, default=14, help="warp instructions per vote")
    a = ap.parse_args()
    cost = np.array([float(x) for x in a.cost.split(",")])
    emu = O.HostEmu()
    h = emu.load(O.scene_path(a.scene))
    cam = O.default_camera_bytes()
    st = O.settings_bytes(sun=bool(a.sun), bounces=a.bounces, max_samples=1 << 30)
    P = lambda x: x.ctypes.data_as(ctypes.c_void_p)
    W, H = 1920, 1080
    rng = np.random.default_rng(1)
    tiles = [(int(rng.integers(0, W // 8)), int(rng.integers(0, H // 4))) for _ in range(a.tiles)]
    policies = [
        ("exclusive argmax", [1, 1, 1, 1, 1], 0.0, 0),
        ("overlap argmax", [1, 1, 1, 1, 1], 0.0, 1),
        ("overlap sticky4", [1, 1, 1, 1, 1], 4.0, 1),
        ("overlap sticky8", [1, 1, 1, 1, 1], 8.0, 1),
        ("overlap T1.3", [1, 1.3, 1, 1, 1], 0.0, 1),
        ("overlap T1.6", [1, 1.6, 1, 1, 1], 0.0, 1),
        ("overlap N1.3", [1.3, 1, 1, 1, 1], 0.0, 1),
        ("overlap T1.3 shade.7", [1, 1.3, .7, .7, .7], 0.0, 1),
        ("overlap T1.3 shade1.3", [1, 1.3, 1.3, 1.3, 1.3], 0.0, 1),
        ("overlap T1.3 sticky4", [1, 1.3, 1, 1, 1], 4.0, 1),
    ]
    print("%-24s %9s %8s %8s  %s" % ("policy", "winst/ray", "simt", "nodes/ry", "lanes per issued step (N T H U E)"))
    policies = [(n, w, s_, o, 1) for n, w, s_, o in policies] + [("pool%d argmax" % k, [1, 1, 1, 1, 1], 0.0, 1, k) for k in (2, 3, 4, 6, 8)] \
        + [("pool%d sticky8" % k, [1, 1, 1, 1, 1], 8.0, 1, k) for k in (2, 4)]
    for name, w, sticky, overlap, pool in policies:
        out = (ctypes.c_ulonglong * 12)()
        wv = np.array(w, np.float32)
        for tx, ty in tiles:
            emu.L.emu_warp_sim(h, P(cam), P(st), W, H, tx, ty, 1, a.spp, P(wv), ctypes.c_float(sticky), overlap, pool, out)
        ex = np.array(out[0:5], float)
        ln = np.array(out[5:10], float)
        rays, nodes = out[10], out[11]
        issued = float(np.dot(ex, cost) + ex.sum() * a.vote)
        useful = float(np.dot(ln, cost)) / 32.0
        print("%-24s %9.1f %8.3f %8.2f  %s" % (name, issued / rays, useful / issued, nodes / rays, np.round(ln / np.maximum(ex, 1), 1)))


if __name__ == "__main__":
    main()
