import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from arrow1_b200 import capi, columns as cols
from arrow1_b200.columns import Column
ctx = capi.Context(0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16_000_000
g = torch.Generator(device="cuda"); g.manual_seed(2)
vals = torch.randint(-2**62, 2**62, (n,), device="cuda", dtype=torch.int64, generator=g)
nb = (n + 7) // 8
vbits = torch.full((nb + 64,), 255, device="cuda", dtype=torch.uint8); vbits[::13] = 0xFE
c = Column(cols.INT64, vals.view(torch.uint8), vbits, n, 0, -1)
out = torch.empty(n, dtype=torch.int64, device="cuda")
scratch = ctx.empty(capi.sort_scratch_bytes(ctx, n))
for _ in range(2):
    capi.sort_indices(ctx, c, 0, 0, out, scratch)
torch.cuda.synchronize()
print("ok")
