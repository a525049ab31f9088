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
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
for dt in ("int64", "int32", "float64"):
    g = torch.Generator(device="cuda"); g.manual_seed(2)
    if dt == "int64":
        vals = torch.randint(-2**62, 2**62, (n,), device="cuda", dtype=torch.int64, generator=g); t = cols.INT64
    elif dt == "int32":
        vals = torch.randint(-2**31, 2**31 - 1, (n,), device="cuda", dtype=torch.int32, generator=g); t = cols.INT32
    else:
        vals = torch.rand((n,), device="cuda", dtype=torch.float64, generator=g) * 1e14; t = cols.DOUBLE
    nb = (n + 7) // 8
    vbits = torch.full((nb + 64,), 255, device="cuda", dtype=torch.uint8); vbits[::13] = 0xFE
    c = Column(t, vals.view(torch.uint8), vbits, n, 0, -1)
    out = torch.empty(n, dtype=torch.int64, device="cuda")
    scratch = ctx.empty(capi.sort_scratch_bytes(ctx, n))
    for _ in range(2): capi.sort_indices(ctx, c, 0, 0, out, scratch)
    L.b200_ctx_set_kernel_timing(ctx._h, 1)
    iters = 3
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): capi.sort_indices(ctx, c, 0, 0, out, scratch)
    e1.record(); torch.cuda.synchronize()
    h, hn = ktime("sort_hist"); p, pn = ktime("sort_pass")
    print("%s n=%d total %.3f ms/sort; hist %.3f ms (x%d); pass total %.3f ms over %d launches" % (
        dt, n, e0.elapsed_time(e1) / iters, h / hn, hn, p / iters, pn // iters))
    L.b200_ctx_set_kernel_timing(ctx._h, 0)
    del scratch
