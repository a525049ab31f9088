#!/usr/bin/env python
"""Single-GPU probe of the range-partition kernels (count / scatter) at 100 M rows: kernel times, for ncu."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from arrow1_b200 import capi, compute, synth
import bench

n = int(os.environ.get("ROWS", "100000000"))
parts = int(os.environ.get("PARTS", "3"))
ctx = compute.default_context(0)
wl = bench.make_workload("sort_i64", n)
col = wl.make(synth.Gen("cuda:0"))[0]
spl = [((1 << 64) * (q + 1)) // (parts - 1) - 1 for q in range(parts - 2)] + [(1 << 64) - 2]
kb = [capi.PeerBuffer(ctx, n * 8 + 64) for _ in range(parts)]
rb = [capi.PeerBuffer(ctx, n * 4 + 64) for _ in range(parts)]
for it in range(4):
    if it == 1:
        capi.lib().b200_ctx_set_kernel_timing(ctx._h, 1)
    plan = capi.RangePartitionPush(ctx, col, 0, spl, parts)
    plan.push(0, [b.ptr for b in kb], [b.ptr for b in rb], [0] * parts, rows32=True)
torch.cuda.synchronize()
print("counts", plan.counts)
print({k: v for k, v in bench.kernel_times(ctx, bench.ALL_KERNELS).items() if v[1]})
