#!/usr/bin/env python
"""Per-phase device times of dist_sort_indices (torchrun, one rank per GPU):
   python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 scripts/dist_sort_breakdown.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda:%d" % local))
from arrow1_b200 import compute, synth, dist as bd
import bench

n = int(os.environ.get("ROWS", "100000000"))
ctx = compute.default_context(local)
wl = bench.make_workload("sort_i64", n)
cols = wl.make(synth.Gen("cuda:%d" % local), start=rank * n)
ops = bd.CudaOps(ctx)
for _ in range(3):
    bd.dist_sort_indices(ops, cols[0], 0, rank * n)
acc = {}
K = 5
for _ in range(K):
    dist.barrier()
    torch.cuda.synchronize()
    bd.TRACE = []
    bd.dist_sort_indices(ops, cols[0], 0, rank * n)
    torch.cuda.synchronize()
    tr = bd.TRACE
    for (a, ea), (b, eb) in zip(tr[:-1], tr[1:]):
        acc[b] = acc.get(b, 0.0) + ea.elapsed_time(eb)
    acc["total"] = acc.get("total", 0.0) + tr[0][1].elapsed_time(tr[-1][1])
bd.TRACE = None
# per-kernel device times of one more round (b200_ctx_set_kernel_timing)
from arrow1_b200 import capi
capi.lib().b200_ctx_set_kernel_timing(ctx._h, 1)
for _ in range(K):
    bd.dist_sort_indices(ops, cols[0], 0, rank * n)
torch.cuda.synchronize()
kt = bench.kernel_times(ctx, bench.ALL_KERNELS)
capi.lib().b200_ctx_set_kernel_timing(ctx._h, 0)
acc["| kernels:"] = 0.0
for k, (ms, cnt) in kt.items():
    if cnt:
        acc["%s(x%d)" % (k, cnt // K)] = ms
for r in range(world):
    dist.barrier()
    if r == rank:
        print("rank %d: " % rank + "  ".join("%s %.2f" % (k, v / K) for k, v in acc.items()), flush=True)
dist.destroy_process_group()
