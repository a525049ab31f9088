"""Quick on-GPU timing probe (not the bench): filter and sort kernels at a few sizes."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from arrow1_b200 import capi, columns as cols
from arrow1_b200.columns import Column

ctx = capi.Context(0)
def timeit(fn, iters=10, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters

what = sys.argv[1:] or ["filter", "sort"]
if "filter" in what:
    for n in (10_000_000, 100_000_000):
        g = torch.Generator(device="cuda"); g.manual_seed(1)
        vals = torch.randint(-2**62, 2**62, (n,), device="cuda", dtype=torch.int64, generator=g)
        nb = (n + 7) // 8
        mbits = torch.randint(0, 256, (nb + 64,), device="cuda", dtype=torch.uint8, generator=g)
        vbits = torch.full((nb + 64,), 255, device="cuda", dtype=torch.uint8)
        vbits[::13] = 0xFE
        v = Column(cols.INT64, vals.view(torch.uint8), vbits, n, 0, -1)
        m = Column(cols.BOOL, mbits, None, n, 0, 0)
        plan = capi.FilterPlan(ctx, m, 0, sync=True)
        out_len = plan.out_length
        od = ctx.empty(n * 8); ov = ctx.empty(nb)
        t_plan = timeit(lambda: capi.FilterPlan(ctx, m, 0, sync=False))
        t_apply = timeit(lambda: plan.apply(v, od, ov))
        bytes_ = n * 8.25 + out_len * 8.125
        print("filter n=%d sel=%.3f plan %.3f ms apply %.3f ms -> apply %.1f GB/s algorithmic, %.2f Grows/s" % (
            n, out_len / n, t_plan, t_apply, bytes_ / t_apply / 1e6, n / t_apply / 1e6))
if "sort" in what:
    for n in (10_000_000, 100_000_000):
        for dt in ("int64", "float64", "int32"):
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
            ms = timeit(lambda: capi.sort_indices(ctx, c, 0, 0, out, scratch), iters=5, warm=2)
            print("sort %s n=%d: %.3f ms -> %.2f Grows/s" % (dt, n, ms, n / ms / 1e6))
            del scratch
