#!/usr/bin/env python3
"""Per-source-line summary of an ncu report (SASS page joined with nvdisasm line info of the SAME build).

    python tools/ncu_lines.py gpurun_out/render_full.ncu-rep [kernel-substring] [--top N]
        [--regions "drt_phased.cuh:500-506=vote,154-226=node,..."] > profiles/rNN_lines.txt

--regions sums program-order regions: an instruction inlined from another file (math, rng, intrinsics) is counted in
the region of the last instruction before it (in SASS order) that maps to a listed line range of the region file.

ncu's own CUDA-source page needs the source tree at the path the box compiled from; the GPU box has a scratch
path, so we join by instruction offset instead: `ncu --page source --csv` gives per-SASS-instruction counters in
program order, `nvdisasm -g` of the cubin inside libdrt.so gives file:line per instruction offset.
"""
import csv
import collections
import io
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "dustraytracer_b200", "libdrt.so")


def sass_lines(kernel_substr):
    """[(offset, file, line, text)] for the first function whose mangled name contains kernel_substr."""
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", LIB], cwd=tmp, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    out = []
    for f in sorted(os.listdir(tmp)):
        if "sm_100a" not in f:
            continue
        txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, f)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
        cur, active, fl = None, False, ("?", 0)
        for ln in txt.splitlines():
            m = re.match(r"^\.text\.(\S+):", ln)
            if m:
                cur = m.group(1)
                active = kernel_substr in cur and not out
                continue
            if not active:
                continue
            m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
            if m:
                fl = (os.path.basename(m.group(1)), int(m.group(2)))
                continue
            m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
            if m:
                out.append((int(m.group(1), 16), fl[0], fl[1], m.group(2).strip()))
        if out:
            break
    return out


def main():
    rep = sys.argv[1]
    ksub = sys.argv[2] if len(sys.argv) > 2 and not sys.argv[2].startswith("--") else "k_renderILb0ELb0ELi0E"
    top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 45
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hdr_i]
    col = {n: i for i, n in enumerate(hdr)}
    body = [r for r in rows[hdr_i + 1:] if r and r[0].startswith("0x")]
    base = int(body[0][0], 16)
    sass = {o: (f, l, t) for o, f, l, t in sass_lines(ksub)}
    agg = collections.defaultdict(lambda: [0, 0, 0, 0])  # inst, thread inst, samples, sass count
    regions, rfile = [], None
    if "--regions" in sys.argv:
        spec = sys.argv[sys.argv.index("--regions") + 1]
        rfile, _, rest = spec.partition(":")
        for item in rest.split(","):
            rng, _, name = item.partition("=")
            lo, _, hi = rng.partition("-")
            regions.append((int(lo), int(hi or lo), name))
    ragg = collections.defaultdict(lambda: [0, 0, 0, 0, 0])  # inst, thread inst, samples, sass count, max executions of one instruction
    cur = "(before first region line)"
    tot_i = tot_t = tot_s = 0
    miss = 0
    for r in body:
        off = int(r[0], 16) - base
        f, l, t = sass.get(off, ("?", 0, ""))
        if f == "?":
            miss += 1
        ie, te, sm = int(r[col["Instructions Executed"]]), int(r[col["Thread Instructions Executed"]]), int(r[col["# Samples"]])
        a = agg[(f, l)]
        a[0] += ie; a[1] += te; a[2] += sm; a[3] += 1
        if regions:
            if f == rfile:
                for lo, hi, name in regions:
                    if lo <= l <= hi:
                        cur = name
                        break
            g = ragg[cur]
            g[0] += ie; g[1] += te; g[2] += sm; g[3] += 1; g[4] = max(g[4], ie)
        tot_i += ie; tot_t += te; tot_s += sm
    print("kernel %s: %d SASS instructions, %d unmatched; warp-inst executed %d, thread-inst %d (avg %.2f active lanes), samples %d"
          % (ksub, len(body), miss, tot_i, tot_t, tot_t / max(tot_i, 1), tot_s))
    byfile = collections.defaultdict(lambda: [0, 0, 0])
    for (f, l), a in agg.items():
        b = byfile[f]
        b[0] += a[0]; b[1] += a[1]; b[2] += a[2]
    print("\nper file:   %-22s %8s %8s %8s" % ("file", "inst%", "lanes", "stall%"))
    for f, b in sorted(byfile.items(), key=lambda kv: -kv[1][0]):
        print("            %-22s %7.2f%% %8.2f %7.2f%%" % (f, 100.0 * b[0] / tot_i, b[1] / max(b[0], 1), 100.0 * b[2] / max(tot_s, 1)))
    if regions:
        print("\nprogram-order regions of %s (inlined callees follow the caller's region):" % rfile)
        print("  %-28s %5s %8s %8s %8s %12s %10s" % ("region", "sass", "inst%", "lanes", "stall%", "max issues", "inst/issue"))
        for name, g in sorted(ragg.items(), key=lambda kv: -kv[1][0]):
            print("  %-28s %5d %7.2f%% %8.2f %7.2f%% %12d %10.1f" % (name, g[3], 100.0 * g[0] / tot_i, g[1] / max(g[0], 1),
                                                                100.0 * g[2] / max(tot_s, 1), g[4], g[0] / max(g[4], 1)))
    print("\ntop lines by issued warp instructions:")
    print("  %-22s %5s %5s %8s %8s %8s" % ("file", "line", "sass", "inst%", "lanes", "stall%"))
    for (f, l), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print("  %-22s %5d %5d %7.2f%% %8.2f %7.2f%%" % (f, l, a[3], 100.0 * a[0] / tot_i, a[1] / max(a[0], 1), 100.0 * a[2] / max(tot_s, 1)))


if __name__ == "__main__":
    main()
