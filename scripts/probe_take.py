#!/usr/bin/env python
"""Take throughput on one GPU: 100 M int64 values gathered by (a) the sort permutation (random), (b) iota."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from arrow1_b200 import capi, compute, synth
import bench

n = int(os.environ.get("ROWS", "100000000"))
ctx = compute.default_context(0)
col = bench.make_workload("sort_i64", n).make(synth.Gen("cuda:0"))[0]
idx = capi.sort_indices(ctx, col, 0)
iota = torch.arange(n, dtype=torch.int64, device="cuda:0")
for name, ix in (("sort permutation", idx), ("iota", iota)):
    for _ in range(3):
        capi.take(ctx, col, ix, True, want_null_count=False)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        capi.take(ctx, col, ix, False, want_null_count=False)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print("take 100M int64 by %-16s %.3f ms  %.3g rows/s  %.0f GB/s (8 B index + 8 B value in, 8 B + 1 bit out)" % (name, ms, n / ms * 1e3, n * 24.125 / ms / 1e6))
