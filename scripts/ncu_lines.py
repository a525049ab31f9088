#!/usr/bin/env python
"""Per-CUDA-source-line stall samples and executed instructions of one kernel in an .ncu-rep
(needs -lineinfo + --import-source on).  usage: ncu_lines.py report.ncu-rep [min_pct]"""
import csv, io, subprocess, sys
from collections import defaultdict

rep = sys.argv[1]
min_pct = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
agg = defaultdict(lambda: [0, 0, ""])
fname, cur, hdr, kernel_seen = None, None, None, 0
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]
        continue
    if r[0] == "Function Name":
        continue
    if r[0] == "Line No":
        hdr = r
        si, ie = hdr.index("# Samples"), hdr.index("Instructions Executed")
        continue
    if hdr is None:
        continue
    if r[0] != "":
        cur = (fname, int(r[0]))
        agg[cur][2] = r[1].strip()[:90]
    elif cur is not None and len(r) > ie:
        try:
            agg[cur][0] += int(r[si] or 0)
            agg[cur][1] += int(r[ie] or 0)
        except ValueError:
            pass
tot_s = sum(v[0] for v in agg.values()) or 1
tot_i = sum(v[1] for v in agg.values()) or 1
print("total samples %d, warp instructions %d" % (tot_s, tot_i))
for (f, ln), v in sorted(agg.items(), key=lambda kv: (kv[0][0], kv[0][1])):
    if 100.0 * v[0] / tot_s >= min_pct or 100.0 * v[1] / tot_i >= min_pct:
        print("%-12s %4d  samples %5.1f%%  instr %5.1f%%  %s" % (f, ln, 100.0 * v[0] / tot_s, 100.0 * v[1] / tot_i, v[2]))
