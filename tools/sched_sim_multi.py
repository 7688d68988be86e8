#!/usr/bin/env python3
"""Design tool: warp-schedule simulator for lanes that own J pixels each (multi-context kernel). CPU only.
tests/hostemu emu_warp_sim_multi runs one warp with the REAL step functions; prints lanes per issue and modelled
warp instructions per ray: cost = N*(cN+oN) + T*(cT+oT) + S*(cS+oS) + draws*cD + votes*cV + ballots*cB.
    python tools/sched_sim_multi.py [--scene room] [--tiles 24] [--spp 8]"""
import argparse, ctypes, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracles as O  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default="room")
    ap.add_argument("--tiles", type=int, default=24)
    ap.add_argument("--spp", type=int, default=8)
    ap.add_argument("--bounces", type=int, default=8)
    ap.add_argument("--cost", default="132,71,70,15,16,10")
    ap.add_argument("--over", default="23,18,40", help="extra instructions per NODE / TRI / SHADE issue when J > 1 (context select, state round trip)")
    a = ap.parse_args()
    cN, cT, cS, cD, cV, cB = [float(x) for x in a.cost.split(",")]
    oN, oT, oS = [float(x) for x in a.over.split(",")]
    emu = O.HostEmu()
    emu.L.emu_warp_sim_multi.restype = None
    h = emu.load(O.scene_path(a.scene))
    cam = O.default_camera_bytes()
    st = O.settings_bytes(sun=True, bounces=a.bounces, max_samples=1 << 30)
    W, H = 1920, 1080
    rng = np.random.RandomState(1)
    tiles = [(int(rng.randint(0, W // 8 - 4)), int(rng.randint(0, H // 4))) for _ in range(a.tiles)]
    print("%-34s %6s %6s %6s %8s %8s" % ("policy", "N", "T", "S", "lanes", "inst/ray"))
    for J in (1, 2, 3, 4):
        for smin in (0, 12, 20, 26):
            for stick in (6,):
                pol = np.array([1, 1, 1, smin, 0, 0, stick, stick, stick, 0, 0, 0], np.float32)
                out = np.zeros(16, np.uint64)
                for (tx, ty) in tiles:
                    # J tiles side by side; every J-th tile column so that the same pixels are covered for every J
                    emu.L.emu_warp_sim_multi(h, O.P(cam), O.P(st), W, H, tx, ty, 1, a.spp, J, O.P(pol), O.P(out))
                o = out.astype(np.float64)
                rays = o[8]
                ov = (0, 0, 0) if J == 1 else (oN, oT, oS)
                cost = o[0] * (cN + ov[0]) + o[1] * (cT + ov[1]) + o[2] * (cS + ov[2]) + o[9] * cD + o[6] * cV + o[7] * cB
                lanes = [o[3 + k] / max(o[k], 1) for k in range(3)]
                allanes = (o[3] * cN + o[4] * cT + o[5] * cS) / max(o[0] * cN + o[1] * cT + o[2] * cS, 1)
                print("J=%d shade>=%-2d stick %-2d            %6.1f %6.1f %6.1f %8.1f %8.1f" % (J, smin, stick, lanes[0], lanes[1], lanes[2], allanes, cost / rays))


if __name__ == "__main__":
    main()
