#!/usr/bin/env python
"""Generates tests/golden/q1_sf100.json: the reference's answer to config C5 (TPC-H Q1-shaped pipeline over the
600 M synthetic lineitem rows of arrow1_b200/synth.py) — run HERE, where /root/reference is compiled into oracle/_ref.

  mask        = shipdate <= 10471                      scalar_compare.cc semantics (C++ <= on int32), stated with numpy
  Filter      = compute::Filter (oracle/_ref) of returnflag, linestatus, quantity, extendedprice, discount, shard by shard
  disc_price  = extendedprice * (1 - discount)         two IEEE double operations, like multiply / subtract
  GroupBy     = compute::internal::GroupBy over ChunkedArrays (chunks of <= 2^27 rows), hash_sum of quantity,
                extendedprice, disc_price, discount and of an int64 column of ones (= the counts; multi-batch hash_count
                is broken upstream).  Sums therefore accumulate in ROW ORDER over all ~578 M rows of a group
                set, exactly as the reference's single-threaded GroupBy does.
Besides the reference's sums the fixture holds each group's sum in extended precision (numpy longdouble, pairwise), so that
the test can say which side drifts when the 1e-12 tolerance is tight.  Sums are stored as float.hex().
usage: python scripts/make_q1_fixture.py [rows]   (default 600000000; ~25 GB of host memory, a few minutes)"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402

import oracle  # noqa: E402
from arrow1_b200 import compute, synth  # noqa: E402
from arrow1_b200.columns import Column  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 600_000_000
SHARD = 1 << 27
assert oracle.have_reference(), "needs oracle/_ref (the compiled reference)"
gen = synth.Gen(None)
chunks = {k: [] for k in ("rf", "ls", "qty", "price", "disc", "dprice", "ones")}
t0 = time.time()
for lo in range(0, N, SHARD):
    n = min(SHARD, N - lo)
    ship, rf, ls, qty, price, disc = gen.q1_lineitem(n, start=lo)
    mask = Column.from_numpy(ship.values() <= compute.Q1_SHIPDATE_CUTOFF)
    f = [oracle.filter(c, mask, 0, kind="reference").values().copy() for c in (rf, ls, qty, price, disc)]
    chunks["rf"].append(f[0])
    chunks["ls"].append(f[1])
    chunks["qty"].append(f[2])
    chunks["price"].append(f[3])
    chunks["disc"].append(f[4])
    chunks["dprice"].append(f[3] * (1.0 - f[4]))
    chunks["ones"].append(np.ones(len(f[0]), dtype=np.int64))
    print("shard at %d: %d rows kept, %.0f s" % (lo, len(f[0]), time.time() - t0), flush=True)
    del ship, rf, ls, qty, price, disc, mask, f
cols = {k: [Column.from_numpy(a) for a in v] for k, v in chunks.items()}
G, aggs, keys = oracle.groupby_sum_chunked([cols["rf"], cols["ls"]],
                                           [cols["qty"], cols["price"], cols["dprice"], cols["disc"], cols["ones"]], 16)
print("reference GroupBy: %d groups, %.0f s" % (G, time.time() - t0), flush=True)
rfk, lsk = keys[0][0], keys[1][0]
# extended-precision sums per group
allk = [np.concatenate(chunks["rf"]).astype(np.uint16) | (np.concatenate(chunks["ls"]).astype(np.uint16) << 8)]
exact = {}
for name in ("qty", "price", "dprice", "disc"):
    v = np.concatenate(chunks[name])
    exact[name] = []
    for g in range(G):
        sel = allk[0] == (int(rfk[g]) | (int(lsk[g]) << 8))
        s = v[sel].astype(np.longdouble).sum()
        hi = float(s)
        exact[name].append([hi.hex(), float(s - np.longdouble(hi)).hex()])
    del v
out = {
    "rows": N, "kept_rows": int(sum(len(a) for a in chunks["rf"])), "groups": int(G),
    "returnflag": [int(x) for x in rfk], "linestatus": [int(x) for x in lsk],
    "count": [int(x) for x in aggs[4][0]],
    "reference_sum": {n: [float(x).hex() for x in aggs[i][0]] for i, n in enumerate(("qty", "price", "dprice", "disc"))},
    "extended_sum_hi_lo": exact,
    "how": "scripts/make_q1_fixture.py: oracle/_ref (unmodified reference) Filter + chunked GroupBy(hash_sum) in row order",
}
path = os.path.join(ROOT, "tests", "golden", "q1_sf100.json" if N == 600_000_000 else "q1_%d.json" % N)
json.dump(out, open(path, "w"), indent=1)
print("wrote", path, "%.0f s" % (time.time() - t0))
