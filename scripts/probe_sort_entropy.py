"""Is the onesweep pass bound by the SM side or by DRAM scatter granularity?  Same row count, same
8 live passes, but each digit takes only `k` distinct values, so the digit-contiguous segments a
tile writes are 256/k times longer (fewer, larger DRAM bursts)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as C
import torch
from arrow1_b200 import capi, columns as cols
from arrow1_b200.columns import Column
ctx = capi.Context(0)
L = capi.lib()
def ktime(name):
    ms, n = C.c_double(0), C.c_int64(0)
    L.b200_ctx_kernel_time(ctx._h, name.encode(), C.byref(ms), C.byref(n))
    return ms.value, n.value
n = 100_000_000
g = torch.Generator(device="cuda"); g.manual_seed(2)
base = torch.randint(-2**62, 2**62, (n,), device="cuda", dtype=torch.int64, generator=g)
out = torch.empty(n, dtype=torch.int64, device="cuda")
scratch = ctx.empty(capi.sort_scratch_bytes(ctx, n))
for name, mask in (("256 values/digit", -1), ("64 values/digit", 0x3F3F3F3F3F3F3F3F), ("16 values/digit", 0x0F0F0F0F0F0F0F0F),
                   ("4 values/digit", 0x0303030303030303), ("2 values/digit", 0x0101010101010101)):
    vals = base & mask
    c = Column(cols.UINT64, vals.view(torch.uint8), None, n, 0, 0)
    for _ in range(2): capi.sort_indices(ctx, c, 0, 0, out, scratch)
    L.b200_ctx_set_kernel_timing(ctx._h, 1)
    iters = 3
    for _ in range(iters): capi.sort_indices(ctx, c, 0, 0, out, scratch)
    torch.cuda.synchronize()
    h, hn = ktime("sort_hist"); p, pn = ktime("sort_pass")
    print("%-18s hist %.3f ms; pass avg %.3f ms over %d launches/sort" % (name, h / hn, p / pn, pn // iters))
    L.b200_ctx_set_kernel_timing(ctx._h, 0)
