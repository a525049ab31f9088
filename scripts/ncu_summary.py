#!/usr/bin/env python
"""Summarises an .ncu-rep (read here, no GPU): key metrics per captured launch, stall breakdown,
and the hottest SASS instructions.  usage: ncu_summary.py report.ncu-rep [launch_index]"""
import csv, subprocess, sys, io
from collections import Counter

rep = sys.argv[1]
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "launch__shared_mem_per_block_dynamic", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_sectors_srcunit_tex_op_write.sum",
        "smsp__inst_executed_op_shared_atom.sum", "l1tex__t_bytes_pipe_lsu_mem_local_op_ld.sum", "l1tex__t_bytes_pipe_lsu_mem_local_op_st.sum"]
for w in want:
    if w in hdr:
        i = hdr.index(w)
        print("%-70s %-14s %s" % (w, units[i], [r[i][:60] for r in data]))
print("--- stalls per issue (launch %d)" % which)
st = []
for i, h in enumerate(hdr):
    if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio"):
        st.append((float(data[which][i] or 0), h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]))
for v, n in sorted(st, reverse=True)[:8]:
    print("  %-22s %.2f" % (n, v))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
sec = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"] + [len(rows)]
start, end = sec[which], sec[which + 1]
h = rows[start + 1]
si, so, ie = h.index("# Samples"), h.index("Source"), h.index("Instructions Executed")
d = rows[start + 2:end]
tot = sum(int(r[si] or 0) for r in d)
print("--- %d SASS instructions, %d samples; hottest:" % (len(d), tot))
cols = [c for c in ("stall_short_sb", "stall_long_sb", "stall_barrier", "stall_wait", "stall_lg", "stall_mio", "stall_branch_resolving", "stall_membar") if c in h]
ci = [h.index(c) for c in cols]
print("   idx samples  instr | " + " ".join(c[6:] for c in cols))
for k, r in enumerate(d):
    s = int(r[si] or 0)
    if s > tot * 0.012:
        print("%6d %6d %5.1f%% x%-9s %-70s %s" % (k, s, 100.0 * s / tot, r[ie], r[so][:70], " ".join(r[c] for c in ci)))
