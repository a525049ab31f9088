#!/usr/bin/env python
"""Why is consume_selected slow?  One value column, 8 groups, 200 M rows: time + launches per variant."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from arrow1_b200 import capi, columns as cols, compute, synth
from arrow1_b200.columns import AggSpec

n = 200_000_000
ctx = compute.default_context(0)
ship, rf, ls, qty, price, disc = synth.Gen("cuda:0").q1_lineitem(n, 0)
print("key types", rf.type, ls.type, "validity", rf.validity is not None, ls.validity is not None, "val validity", qty.validity is not None, qty.null_count)
mask = capi.compare_scalar(ctx, ship, "less_equal", compute.Q1_SHIPDATE_CUTOFF)
print("mask validity", mask.validity is not None, mask.null_count, mask.offset)
aggs = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_MEAN)]
def run(name, **kw):
    for _ in range(2):
        capi.group_by(ctx, [rf, ls], aggs, [qty, qty], expected_groups=8, **kw)
    torch.cuda.synchronize()
    l0 = ctx.launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    capi.group_by(ctx, [rf, ls], aggs, [qty, qty], expected_groups=8, **kw)
    e1.record(); torch.cuda.synchronize()
    print("%-28s %.2f ms  launches %d" % (name, e0.elapsed_time(e1), ctx.launches - l0))
run("no selection")
run("selection", selection=mask)
from arrow1_b200.columns import Column
ones = Column(cols.BOOL, torch.full(((n + 31) // 32 * 4 + 8,), 255, dtype=torch.uint8, device="cuda:0"), None, n, 0, 0)
run("all-ones selection", selection=ones)
half = Column(cols.BOOL, torch.randint(0, 256, ((n + 31) // 32 * 4 + 8,), dtype=torch.uint8, device="cuda:0"), None, n, 0, 0)
run("random 50% selection", selection=half)
inv = Column(cols.BOOL, ~mask.data, None, n, 0, 0)
run("inverted real mask (3.6%)", selection=inv)
r = torch.rand(n, device="cuda:0") < 0.964
import numpy as np
w = torch.tensor([1, 2, 4, 8, 16, 32, 64, 128], dtype=torch.uint8, device="cuda:0")
pad = torch.zeros(((n + 31) // 32 * 32 + 64,), dtype=torch.bool, device="cuda:0"); pad[:n] = r
packed = (pad.view(-1, 8).to(torch.uint8) * w).sum(dim=1, dtype=torch.uint8)
rnd = Column(cols.BOOL, packed, None, n, 0, 0)
run("random 96.4% selection", selection=rnd)
import bench
capi.lib().b200_ctx_set_kernel_timing(ctx._h, 1)
capi.group_by(ctx, [rf, ls], aggs, [qty, qty], expected_groups=8, selection=mask)
print(bench.kernel_times(ctx, bench.ALL_KERNELS))
