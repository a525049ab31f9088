"""Hash repartitioning for the multi-GPU high-cardinality group-by, on one
GPU: the rows are partitioned for P destinations by the CUDA kernels, every destination's rows are
aggregated by its own handle from explicit global row numbers (b200_groupby_consume_rows), and the
groups of all destinations, ranked by global first row, must equal the oracle's single-pass output
bit for bit (float sums within 1e-12).  The NCCL all-to-all between the two steps is exercised by
tests/test_dist_cpu.py (gloo) and by `bench.py --workload groupby_c4 --gpus N`."""
import numpy as np
import pytest
import torch

import oracle
from arrow1_b200 import capi, columns as cols
from arrow1_b200.columns import AggSpec, Column
from arrow1_b200.dist import CudaOps
from gen import rand_column
from helpers import GroupRes, abs_sum_scales, assert_group_equal, run_groupby

pytestmark = pytest.mark.gpu
ORACLE = "reference" if oracle.have_reference() else "port"
AGGS = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_COUNT), AggSpec(cols.HASH_MIN_MAX), AggSpec(cols.HASH_MEAN)]


def _repartitioned(ctx, keys, aggs, args, n_parts, row_base=0):
    ops = CudaOps(ctx)
    dkeys = [k.to_device() for k in keys]
    darg = args[0].to_device()
    dargs = [darg] * len(args)
    key_types, arg_types = [k.type for k in keys], [a.type for a in args]
    planner = capi.GroupBy(ctx, key_types, aggs, arg_types, 0)
    columns = dkeys + [darg]
    counts, datas, valids, rows = ops.partition(planner, dkeys, columns, n_parts, row_base)
    planner.close()
    n = keys[0].length
    assert sum(counts) == n
    assert sorted(rows.cpu().tolist()) == list(range(row_base, row_base + n)), "every row lands exactly once"
    merged = []
    seen = set()
    start = 0
    for p in range(n_parts):
        c = counts[p]
        part_cols = []
        for col, d, v in zip(columns, datas, valids):
            w = cols.BITS[col.type] // 8
            part_cols.append(ops.column_from_exchange(col.type, d[start * w:(start + c) * w].clone(),
                                                      None if v is None else v[start:start + c].clone(), c))
        gb = capi.GroupBy(ctx, key_types, aggs, arg_types, 0)
        if c:
            gb.consume_rows(part_cols[:len(keys)], [part_cols[-1]] * len(args), rows[start:start + c].clone())
        out = gb.finalize()
        fr = gb.first_rows().cpu().numpy().view(np.uint64)
        gb.close()
        assert len(fr) == out.num_groups
        assert (np.diff(fr.astype(np.int64)) > 0).all(), "groups come out in first-row order"
        hk = [(k.to_host().values().copy(), k.to_host().valid()) for k in out.keys]
        ha = [[(c_.to_host().values().copy(), c_.to_host().valid()) for c_ in a] for a in out.aggs]
        for g in range(out.num_groups):
            ident = tuple((hk[i][0][g].tobytes() if hk[i][1][g] else None) for i in range(len(keys)))
            assert ident not in seen, "destinations must own disjoint key sets"
            seen.add(ident)
            merged.append((int(fr[g]), p, g))
        start += c
        merged_part = (hk, ha)
        if p == 0:
            parts = []
        parts.append(merged_part)
    merged.sort()
    G = len(merged)
    keys_out = []
    for i in range(len(keys)):
        vals = np.array([parts[p][0][i][0][g] for _, p, g in merged]) if G else parts[0][0][i][0][:0]
        valid = np.array([parts[p][0][i][1][g] for _, p, g in merged], dtype=bool) if G else np.zeros(0, dtype=bool)
        keys_out.append((vals, valid))
    aggs_out = []
    for j in range(len(aggs)):
        lst = []
        for q in range(len(parts[0][1][j])):
            vals = np.array([parts[p][1][j][q][0][g] for _, p, g in merged]) if G else parts[0][1][j][q][0][:0]
            valid = np.array([parts[p][1][j][q][1][g] for _, p, g in merged], dtype=bool) if G else np.zeros(0, dtype=bool)
            lst.append((vals, valid))
        aggs_out.append(lst)
    return GroupRes(G, aggs_out, keys_out), counts


@pytest.mark.parametrize("n_parts", [2, 3, 8])
@pytest.mark.parametrize("kt,at", [(cols.INT64, cols.DOUBLE), (cols.INT32, cols.INT64), (cols.UINT16, cols.FLOAT)])
def test_repartitioned_groupby_matches_oracle(ctx, n_parts, kt, at):
    rng = np.random.default_rng(900 + 7 * n_parts + kt)
    n = 60000
    key = rand_column(rng, kt, n, 0.02, lo=0, hi=20000)
    arg = rand_column(rng, at, n, 0.05, special=True)
    args = [arg] * len(AGGS)
    got, counts = _repartitioned(ctx, [key], AGGS, args, n_parts, row_base=12345)
    want = run_groupby(ORACLE, [key], AGGS, args)
    assert_group_equal(got, want, AGGS, rtol=1e-12, scales=abs_sum_scales([key], AGGS, args))
    assert min(counts) > n // n_parts // 2, "the key hash must balance the destinations: %r" % (counts,)


def test_repartition_multi_key_and_int64_sentinels(ctx):
    rng = np.random.default_rng(77)
    n = 30000
    k0 = rand_column(rng, cols.INT32, n, 0.05, lo=-50, hi=50)
    k1 = rand_column(rng, cols.UINT8, n, 0.05, lo=0, hi=7)
    arg = rand_column(rng, cols.DOUBLE, n, 0.1)
    aggs = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_COUNT)]
    got, _ = _repartitioned(ctx, [k0, k1], aggs, [arg, arg], 4)
    want = run_groupby(ORACLE, [k0, k1], aggs, [arg, arg])
    assert_group_equal(got, want, aggs, rtol=1e-12, scales=abs_sum_scales([k0, k1], aggs, [arg, arg]))
    # single 64-bit key column: the null key and the all-ones key have one fixed owner
    v = rng.integers(-5, 5, size=n).astype(np.int64)
    key = Column.from_numpy(v, rng.random(n) > 0.1)
    got, _ = _repartitioned(ctx, [key], aggs, [arg, arg], 3)
    want = run_groupby(ORACLE, [key], aggs, [arg, arg])
    assert_group_equal(got, want, aggs, rtol=1e-12, scales=abs_sum_scales([key], aggs, [arg, arg]))


def test_pack_validity_round_trip(ctx):
    rng = np.random.default_rng(3)
    for n in (1, 31, 32, 33, 1000, 100003):
        valid = rng.random(n) > 0.3
        t = torch.from_numpy(valid.astype(np.uint8)).to("cuda")
        bitmap, nulls = capi.pack_validity(ctx, t)
        assert nulls == int((~valid).sum())
        got = np.unpackbits(bitmap.cpu().numpy()[: (n + 7) // 8], bitorder="little")[:n].astype(bool)
        np.testing.assert_array_equal(got, valid)
