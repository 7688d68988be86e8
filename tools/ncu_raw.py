#!/usr/bin/env python3
"""Selected raw metrics of an ncu report as `name [unit] = value` lines (the format bench.py's ncu_capture_summary reads).
    python tools/ncu_raw.py gpurun_out/render_full.ncu-rep "header comment" > profiles/rNN_render_phased_raw.txt"""
import csv
import io
import re
import subprocess
import sys

KEEP = re.compile(r"^(Kernel Name|dram__|gpu__time_duration|gpc__cycles_elapsed\.max|l1tex__t_sector_hit_rate|l1tex__data_pipe_lsu_wavefronts\.|"
                  r"l1tex__t_sectors_pipe_lsu_mem_(global|local)_op_(ld|st)\.sum$|lts__t_sector_hit_rate|lts__t_bytes\.sum|launch__|"
                  r"sm__inst_executed|sm__pipe_|sm__throughput|sm__warps_active|smsp__warps_|smsp__issue_active|smsp__inst_executed\.(sum|avg)|"
                  r"smsp__thread_inst_executed|smsp__average_warp|smsp__cycles_active\.avg|sm__cycles_elapsed\.(avg|max)$)")


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    if len(sys.argv) > 2:
        print("# " + sys.argv[2])
    for n, u, v in zip(hdr, units, vals):
        if KEEP.match(n):
            print("%s [%s] = %s" % (n, u, v.replace(",", "")))


if __name__ == "__main__":
    main()
