#!/usr/bin/env python
"""One step of every hot-path workload between cudaProfilerStart / Stop, after an unprofiled warm-up step each, so that
ONE `ncu --profile-from-start off` command captures one launch sequence of each kernel (scripts/gpu_profile.sh).
Without ncu it just runs the steps and prints their wall times (the command must exit 0 on its own first)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402
from arrow1_b200 import capi, synth  # noqa: E402

WORKLOADS = [("sort_i64", 100_000_000), ("filter", 100_000_000), ("groupby_c3", 200_000_000), ("groupby_c4", 200_000_000),
             ("q1_fused", 200_000_000)]
only = sys.argv[1].split(",") if len(sys.argv) > 1 else None
ctx = capi.Context(0)
gen = synth.Gen("cuda:0")
prepared = []
for name, rows in WORKLOADS:
    if only and name not in only:
        continue
    wl = bench.make_workload(name, rows)
    if name == "groupby_c4":
        wl.n_keys = rows // 10
    cols_ = wl.make(gen)
    bufs = wl.prepare(ctx, cols_)
    wl.step(ctx, cols_, bufs)          # warm-up: allocations, function attributes
    torch.cuda.synchronize()
    prepared.append((wl, cols_, bufs))
torch.cuda.profiler.start()
for wl, cols_, bufs in prepared:
    t0 = time.time()
    wl.step(ctx, cols_, bufs)
    torch.cuda.synchronize()
    print("%-12s %8.2f ms (%d rows)" % (wl.name, (time.time() - t0) * 1e3, wl.n), flush=True)
torch.cuda.profiler.stop()
