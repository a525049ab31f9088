#!/usr/bin/env python
"""Offline check of the built library's SASS (no GPU): global loads whose destination registers are read again within
a few instructions -- the pattern that cost gb_bin32_kernel 20 % of its time (a register move placed right behind the
column loads waits for the whole memory latency).  Most hits are ordinary dependent loads (an offset, then its use);
the ones to look at are wide loads of a software-pipelined loop followed by IMAD.MOV / MOV of their registers.
usage: sass_load_use.py [lib.so] [window]"""
import re, subprocess, sys
lib = sys.argv[1] if len(sys.argv) > 1 else "arrow1_b200/libarrow1_b200.so"
win = int(sys.argv[2]) if len(sys.argv) > 2 else 6
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
fn, ins = None, []
def flush():
    if not fn: return
    hits = []
    for i, (addr, text) in enumerate(ins):
        m = re.match(r"(@!?U?P\d+\s+)?LDG\.E(\.\S*)?\s+R(\d+),", text)
        if not m: continue
        op = text.split("R", 1)[0]   # the opcode with its suffixes only (the address has a ".64" of its own)
        width = 4 if ".128" in op else 2 if ".64" in op else 8 if ".256" in op else 1
        base = int(m.group(3)); regs = {"R%d" % (base + k) for k in range(width)}
        for j in range(i + 1, min(i + 1 + win, len(ins))):
            t = ins[j][1]
            if "LDG" in " ".join(t.split()[0:2]): continue
            ops = t.split(None, 1)[1] if " " in t else ""
            dst, _, src = ops.partition(",")
            used = set(re.findall(r"R\d+", src)) | (set(re.findall(r"R\d+", dst)) if t.startswith(("ST", "RED", "ATOM")) else set())
            if used & regs and "BRA" not in t and "EXIT" not in t:
                hits.append((addr, text[:60], ins[j][0], t[:70])); break
    if hits:
        print("== %s: %d load(s) used within %d instructions" % (fn[:90], len(hits), win))
        for h in hits[:6]: print("   %s %-60s -> %s %s" % h)
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        flush(); fn, ins = m.group(1), []; continue
    m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", line)
    if m: ins.append((m.group(1), m.group(2).strip()))
flush()
