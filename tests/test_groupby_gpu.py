"""Grouper + hash aggregate parity: CUDA path (through the C ABI) vs the oracle.  Bit-exact for
keys, group order (first occurrence), counts, min/max (incl. +-0.0 and NaN rules) and integer
sums; float64 sums / means within 1e-12 relative (north_star tolerance)."""
import numpy as np
import pytest
import torch

import oracle
from arrow1_b200 import columns as cols
from arrow1_b200.columns import AggSpec, Column
from conftest import rand_u64
from gen import rand_column
from helpers import abs_sum_scales, assert_group_equal, run_groupby

pytestmark = pytest.mark.gpu
ORACLE = "reference" if oracle.have_reference() else "port"
RTOL = 1e-12

ALL_AGGS = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_COUNT), AggSpec(cols.HASH_COUNT, 1),
            AggSpec(cols.HASH_MIN_MAX), AggSpec(cols.HASH_MIN_MAX, 1), AggSpec(cols.HASH_MEAN)]


@pytest.mark.parametrize("kt", [cols.BOOL, cols.UINT8, cols.INT16, cols.INT32, cols.INT64, cols.UINT64, cols.FLOAT, cols.DOUBLE])
@pytest.mark.parametrize("at", [cols.INT8, cols.UINT16, cols.INT32, cols.UINT64, cols.INT64, cols.FLOAT, cols.DOUBLE])
def test_groupby_random(ctx, kt, at):
    # hash_aggregate_test.cc:704-723 shape (keys in [0,100], lengths 2^10..2^15, null_prob ladder)
    rng = np.random.default_rng(500 + 31 * kt + at)
    for n in (1, 1 << 10, 1 << 15):
        for null_prob in (0.0, 0.01, 0.5, 1.0):
            key = rand_column(rng, kt, n, 0.05, lo=0, hi=100, special=(kt in (cols.FLOAT, cols.DOUBLE)))
            arg = rand_column(rng, at, n, null_prob, special=True)
            args = [arg] * len(ALL_AGGS)
            a = run_groupby("gpu", [key], ALL_AGGS, args, ctx, want_ids=True)
            b = run_groupby(ORACLE, [key], ALL_AGGS, args, want_ids=True)
            assert_group_equal(a, b, ALL_AGGS, rtol=RTOL, scales=abs_sum_scales([key], ALL_AGGS, args))


def test_groupby_empty(ctx):
    key = Column.from_numpy(np.array([], dtype=np.int32))
    arg = Column.from_numpy(np.array([], dtype=np.float64))
    r = run_groupby("gpu", [key], [AggSpec(cols.HASH_SUM)], [arg], ctx)
    assert r.num_groups == 0


def test_groupby_multi_key(ctx):
    rng = np.random.default_rng(11)
    n = 50000
    k0 = rand_column(rng, cols.UINT8, n, 0.1, lo=0, hi=3)
    k1 = rand_column(rng, cols.INT16, n, 0.1, lo=-2, hi=2)
    k2 = rand_column(rng, cols.BOOL, n, 0.1)
    arg = rand_column(rng, cols.DOUBLE, n, 0.2)
    aggs = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_MIN_MAX), AggSpec(cols.HASH_COUNT)]
    a = run_groupby("gpu", [k0, k1, k2], aggs, [arg] * 3, ctx, want_ids=True)
    b = run_groupby(ORACLE, [k0, k1, k2], aggs, [arg] * 3, want_ids=True)
    assert_group_equal(a, b, aggs, rtol=RTOL)


def test_groupby_int64_key_sentinel_and_null(ctx):
    """A single 64-bit key leaves no spare bit: null and the all-ones key live in special slots."""
    v = np.array([-1, 5, -1, 0, 7, 5, -1, 0], dtype=np.int64)  # -1 == the EMPTY sentinel bit pattern
    valid = np.array([1, 1, 1, 0, 1, 1, 0, 1], dtype=bool)
    key = Column.from_numpy(v, valid)
    arg = Column.from_numpy(np.arange(8, dtype=np.float64))
    aggs = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_COUNT)]
    a = run_groupby("gpu", [key], aggs, [arg, arg], ctx, want_ids=True)
    b = run_groupby(ORACLE, [key], aggs, [arg, arg], want_ids=True)
    assert_group_equal(a, b, aggs, rtol=RTOL)
    ukey = Column.from_numpy(v.view(np.uint64), valid)
    a = run_groupby("gpu", [ukey], aggs, [arg, arg], ctx, want_ids=True)
    b = run_groupby(ORACLE, [ukey], aggs, [arg, arg], want_ids=True)
    assert_group_equal(a, b, aggs, rtol=RTOL)


def test_groupby_min_max_zero_sign_and_nan(ctx):
    # std::max/min tie rule: the earlier of +-0.0 wins; NaN never wins but sets has_values
    # (hash_aggregate.cc:663-677)
    vals = np.array([0.0, -0.0, -0.0, 0.0, np.nan, np.nan, 1.0, -0.0, -1.0, 0.0, np.nan, 3.0], dtype=np.float64)
    keys = np.array([0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5], dtype=np.int32)
    for dt, t in ((np.float64, cols.DOUBLE), (np.float32, cols.FLOAT)):
        key = Column.from_numpy(keys)
        arg = Column.from_numpy(vals.astype(dt))
        aggs = [AggSpec(cols.HASH_MIN_MAX)]
        a = run_groupby("gpu", [key], aggs, [arg], ctx)
        b = run_groupby(ORACLE, [key], aggs, [arg])
        assert_group_equal(a, b, aggs)
        # signs compared bitwise by assert_group_equal (view as bytes)


def test_groupby_high_cardinality_growth(ctx):
    """No hint, far more groups than the initial table: rows get deferred, the table grows, parity holds."""
    rng = np.random.default_rng(21)
    n = 400000
    key = Column.from_numpy(rng.integers(0, 1 << 40, size=n, dtype=np.int64))
    arg = Column.from_numpy(rng.random(n))
    aggs = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_MEAN)]
    a = run_groupby("gpu", [key], aggs, [arg, arg], ctx, expected_groups=0)
    b = run_groupby(ORACLE, [key], aggs, [arg, arg])
    assert_group_equal(a, b, aggs, rtol=RTOL)
    a = run_groupby("gpu", [key], aggs, [arg, arg], ctx, expected_groups=n)
    assert_group_equal(a, b, aggs, rtol=RTOL)


def test_groupby_fast_path_spill_and_deferral(ctx):
    """One value column, no hint, ~1 M distinct keys over 2 M rows: every CTA of the fast kernel fills
    its 1024-group table (rows beyond are deferred to the generic kernel) and the CTAs' flushes
    overflow the hint-sized global table (groups go through the spill list and are imported)."""
    rng = np.random.default_rng(77)
    n = 2_000_000
    key = Column.from_numpy(rng.integers(0, 1_000_000, size=n).astype(np.int32))
    arg = Column.from_numpy(rng.integers(-1000, 1000, size=n).astype(np.int64), rng.random(n) > 0.01)
    aggs = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_COUNT), AggSpec(cols.HASH_MIN_MAX)]
    a = run_groupby("gpu", [key], aggs, [arg] * 3, ctx, expected_groups=0)
    b = run_groupby(ORACLE, [key], aggs, [arg] * 3)
    assert_group_equal(a, b, aggs, rtol=RTOL)


def test_groupby_streaming_batches_match_single_batch(ctx):
    """GroupBy's batch loop (hash_aggregate.cc:928-943): consuming row ranges one after the other,
    sliced (offset != 0) — the case the reference itself gets wrong — equals one batch."""
    from arrow1_b200 import capi
    rng = np.random.default_rng(22)
    n = 100000
    kv = rng.integers(0, 300, size=n, dtype=np.int32)
    av = rng.random(n) * 1e6
    avalid = rng.random(n) > 0.1
    key, arg = Column.from_numpy(kv), Column.from_numpy(av, avalid)
    aggs = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_COUNT), AggSpec(cols.HASH_MIN_MAX)]
    gb = capi.GroupBy(ctx, [cols.INT32], aggs, [cols.DOUBLE] * 3)
    dkey, darg = key.to_device(), arg.to_device()
    step = 33333
    for lo in range(0, n, step):
        ln = min(step, n - lo)
        ks = Column(dkey.type, dkey.data, dkey.validity, ln, lo, 0)
        as_ = Column(darg.type, darg.data, darg.validity, ln, lo, -1)
        gb.consume([ks], [as_] * 3, row_base=lo)
    out = gb.finalize()
    gb.close()
    b = run_groupby(ORACLE, [key], aggs, [arg] * 3)
    assert out.num_groups == b.num_groups
    np.testing.assert_array_equal(out.keys[0].to_host().values(), b.keys[0][0])
    np.testing.assert_array_equal(out.aggs[1][0].to_host().values(), b.aggs[1][0][0])
    np.testing.assert_allclose(out.aggs[0][0].to_host().values(), b.aggs[0][0][0], rtol=RTOL)
    np.testing.assert_array_equal(out.aggs[2][0].to_host().values(), b.aggs[2][0][0])
    np.testing.assert_array_equal(out.aggs[2][1].to_host().values(), b.aggs[2][1][0])


def test_groupby_merge_and_export_import(ctx):
    """HashAggregateKernel::merge (NotImplemented in the reference, hash_aggregate.cc:795-798) ==
    what a single table over all rows gives; also through the flat export/import the multi-GPU
    all-gather uses."""
    from arrow1_b200 import capi
    rng = np.random.default_rng(23)
    n = 60000
    kv = rng.integers(0, 5000, size=n, dtype=np.int64)
    av = rng.integers(-1000, 1000, size=n, dtype=np.int32)
    key, arg = Column.from_numpy(kv), Column.from_numpy(av)
    aggs = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_MIN_MAX), AggSpec(cols.HASH_MEAN)]
    b = run_groupby(ORACLE, [key], aggs, [arg] * 3)
    dkey, darg = key.to_device(), arg.to_device()
    half = n // 2
    parts = []
    for lo, ln in ((0, half), (half, n - half)):
        g = capi.GroupBy(ctx, [cols.INT64], aggs, [cols.INT32] * 3)
        g.consume([Column(dkey.type, dkey.data, None, ln, lo, 0)], [Column(darg.type, darg.data, None, ln, lo, 0)] * 3,
                  row_base=lo)
        parts.append(g)
    # (a) direct merge
    parts[0].merge(parts[1])
    out = parts[0].finalize()
    # (b) export -> import into a fresh handle
    g2 = capi.GroupBy(ctx, [cols.INT64], aggs, [cols.INT32] * 3)
    g2.import_state(parts[0].export_state())
    out2 = g2.finalize()
    for o in (out, out2):
        assert o.num_groups == b.num_groups
        np.testing.assert_array_equal(o.keys[0].to_host().values(), b.keys[0][0])
        np.testing.assert_array_equal(o.aggs[0][0].to_host().values(), b.aggs[0][0][0])
        np.testing.assert_array_equal(o.aggs[1][0].to_host().values(), b.aggs[1][0][0])
        np.testing.assert_array_equal(o.aggs[1][1].to_host().values(), b.aggs[1][1][0])
        np.testing.assert_allclose(o.aggs[2][0].to_host().values(), b.aggs[2][0][0], rtol=RTOL)
    for g in parts + [g2]:
        g.close()


def test_hash_agg_kernel_abi_with_dense_ids(ctx):
    """The reference's HashAggregateKernel signature (arg, uint32 ids, num_groups)
    (hash_aggregate.cc:783-789): Grouper ids -> per-aggregate consume -> finalize."""
    from arrow1_b200 import capi
    rng = np.random.default_rng(24)
    n = 20000
    key = rand_column(rng, cols.INT32, n, 0.05, lo=0, hi=50)
    arg = rand_column(rng, cols.DOUBLE, n, 0.1)
    aggs = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_COUNT), AggSpec(cols.HASH_MIN_MAX)]
    b = run_groupby(ORACLE, [key], aggs, [arg] * 3, want_ids=True)
    grouper = capi.GroupBy(ctx, [cols.INT32], [], [])
    ids = grouper.grouper_consume([key.to_device()])
    G = grouper.num_groups()
    np.testing.assert_array_equal(ids.cpu().numpy().view(np.uint32), b.ids)
    assert G == b.num_groups
    uniques = grouper.finalize().keys[0].to_host()
    np.testing.assert_array_equal(uniques.valid(), b.keys[0][1])
    np.testing.assert_array_equal(uniques.values()[uniques.valid()], b.keys[0][0][b.keys[0][1]])
    kernels = capi.GroupBy(ctx, [cols.INT32], aggs, [cols.DOUBLE] * 3)
    darg = arg.to_device()
    for j in range(3):
        kernels.agg_consume(j, darg, ids, G)
    out = kernels.finalize()
    scale = abs_sum_scales([key], aggs, [arg] * 3)[0]
    sv = b.aggs[0][0][1]
    assert (np.abs(out.aggs[0][0].to_host().values()[sv] - b.aggs[0][0][0][sv]) <= RTOL * scale[sv]).all()
    np.testing.assert_array_equal(out.aggs[0][0].to_host().valid(), sv)
    np.testing.assert_array_equal(out.aggs[1][0].to_host().values(), b.aggs[1][0][0])
    v = b.aggs[2][0][1]
    np.testing.assert_array_equal(out.aggs[2][0].to_host().valid(), v)
    np.testing.assert_array_equal(out.aggs[2][0].to_host().values()[v], b.aggs[2][0][0][v])
    np.testing.assert_array_equal(out.aggs[2][1].to_host().values()[v], b.aggs[2][1][0][v])
    grouper.close()
    kernels.close()


def test_groupby_unsupported(ctx):
    from arrow1_b200 import capi
    with pytest.raises(capi.B200Error) as ei:  # summing booleans (hash_aggregate.cc:550-556)
        capi.GroupBy(ctx, [cols.INT32], [AggSpec(cols.HASH_SUM)], [cols.BOOL])
    assert ei.value.code == 10
    with pytest.raises(capi.B200Error) as ei:  # more key columns than the key encoder of this library takes
        capi.GroupBy(ctx, [cols.INT64] * 17, [], [])
    assert ei.value.code == 10
    gb = capi.GroupBy(ctx, [cols.INT64, cols.INT64], [], [])   # wide keys are grouped through the key -> id table,
    with pytest.raises(capi.B200Error) as ei:                   # but their state image cannot be exported
        gb.export_state()
    assert ei.value.code == 10
    gb.close()


def test_groupby_config3_shape(ctx):
    """BASELINE config 3 shape (int32 key % 1000, float64 value, 1 % nulls; sum+count+min_max) at
    20 M rows vs the oracle; sums within 1e-12, everything else exact."""
    n = 20_000_000
    with np.errstate(over="ignore"):
        kv = (rand_u64(7, n) % np.uint64(1000)).astype(np.int32)
        av = (rand_u64(8, n) >> np.uint64(11)).astype(np.float64) * (2.0 ** -53) * 1e14
        valid = (rand_u64(12, n) % np.uint64(100)) != 0
    key, arg = Column.from_numpy(kv), Column.from_numpy(av, valid)
    aggs = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_COUNT), AggSpec(cols.HASH_MIN_MAX)]
    a = run_groupby("gpu", [key], aggs, [arg] * 3, ctx, expected_groups=1000)
    b = run_groupby(ORACLE, [key], aggs, [arg] * 3)
    assert_group_equal(a, b, aggs, rtol=RTOL)
    # checksum of counts == number of valid rows
    assert int(a.aggs[1][0][0].sum()) == int(valid.sum())


def test_groupby_config4_shape(ctx):
    """BASELINE config 4 shape (int64 keys spread over the full range, ~n/10 distinct; sum+mean) at 4 M rows."""
    n = 4_000_000
    with np.errstate(over="ignore"):
        kv = (rand_u64(9, n) % np.uint64(n // 10))
        from conftest import splitmix64
        kv = splitmix64(kv).view(np.int64)
        av = (rand_u64(10, n) >> np.uint64(11)).astype(np.float64) * (2.0 ** -53) * 1e14
    key, arg = Column.from_numpy(kv), Column.from_numpy(av)
    aggs = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_MEAN)]
    a = run_groupby("gpu", [key], aggs, [arg] * 2, ctx, expected_groups=n // 10)
    b = run_groupby(ORACLE, [key], aggs, [arg] * 2)
    assert_group_equal(a, b, aggs, rtol=RTOL)


def test_groupby_low_card_kernel_spill_without_hint(ctx):
    """Several value columns and no hint take the shared-table kernel; with ~1 M distinct keys its
    CTAs overflow into the global table and their final flush spills: parity must hold."""
    rng = np.random.default_rng(78)
    n = 1_500_000
    key = Column.from_numpy(rng.integers(0, 1_000_000, size=n).astype(np.int32))
    a1 = Column.from_numpy(rng.integers(-1000, 1000, size=n).astype(np.int64))
    a2 = Column.from_numpy(rng.integers(0, 1000, size=n).astype(np.int32), rng.random(n) > 0.1)
    aggs = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_MIN_MAX), AggSpec(cols.HASH_COUNT)]
    a = run_groupby("gpu", [key], aggs, [a1, a2, a2], ctx, expected_groups=0)
    b = run_groupby(ORACLE, [key], aggs, [a1, a2, a2])
    assert_group_equal(a, b, aggs, rtol=RTOL)


@pytest.mark.parametrize("n_keys", [1, 2, 3])
@pytest.mark.parametrize("dt", [np.float64, np.int64, np.uint64])
def test_groupby_tiny_byte_keys_register_accumulators(ctx, n_keys, dt):
    """<= 8 groups over uint8 / int8 key columns and a 64-bit value column (the Q1 shape): gb_tiny8_kernel keeps
    the sums and counts in registers.  Also: a 9th..12th group appearing late (those rows are deferred to the generic
    kernel), value nulls, COUNT_NULL, a group made of -0.0 values only (its sum keeps the sign), int64 wrap-around."""
    rng = np.random.default_rng(31 * n_keys + np.dtype(dt).num)
    n = 300_000
    combos = [(1, 2, 3), (1, 5, 3), (7, 2, 3), (7, 5, 9), (200, 2, 3), (200, 5, 250)][:5]
    pick = rng.integers(0, len(combos), size=n)
    pick[n - 5000:] = rng.integers(0, 12, size=5000) % 12          # late: more than 8 groups in the last CTAs
    extra = [(10 + i, 100 + i, 50 + i) for i in range(12)]
    table = np.array(combos + extra[: 12 - len(combos)], dtype=np.int64)
    kcols = []
    for k in range(n_keys):
        vals = table[pick % len(table), k]
        kcols.append(Column.from_numpy(vals.astype(np.uint8 if k != 1 else np.int8)))
    if dt == np.float64:
        v = rng.random(n) * 1e6 - 5e5
        v[pick == 1] = -0.0                                         # a whole group of negative zeros
    elif dt == np.int64:
        v = rng.integers(-2**62, 2**62, size=n, dtype=np.int64)     # sums wrap
    else:
        v = rng.integers(0, 2**63, size=n, dtype=np.uint64)
    arg = Column.from_numpy(v.astype(dt), rng.random(n) > 0.02)
    aggs = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_MEAN), AggSpec(cols.HASH_COUNT), AggSpec(cols.HASH_COUNT, 1)]
    a = run_groupby("gpu", kcols, aggs, [arg] * 4, ctx, expected_groups=6)
    b = run_groupby(ORACLE, kcols, aggs, [arg] * 4)
    assert_group_equal(a, b, aggs, rtol=RTOL)
    # no nulls at all, one key column, streaming in two batches through the same handle shape
    arg2 = Column.from_numpy(v.astype(dt))
    a2 = run_groupby("gpu", kcols, aggs[:2], [arg2] * 2, ctx, expected_groups=6)
    b2 = run_groupby(ORACLE, kcols, aggs[:2], [arg2] * 2)
    assert_group_equal(a2, b2, aggs[:2], rtol=RTOL)


# ---- high-cardinality path (partition, then aggregate in shared memory): forced at test sizes through the
# handle's options; the results must not depend on the path
HP = {"hp_min_rows": 0, "hp_min_groups": 0}
HP_AGGS = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_COUNT), AggSpec(cols.HASH_COUNT, 1), AggSpec(cols.HASH_MEAN)]


@pytest.mark.parametrize("kt", [cols.UINT8, cols.INT32, cols.INT64, cols.UINT64, cols.DOUBLE])
@pytest.mark.parametrize("at", [cols.INT8, cols.INT32, cols.UINT64, cols.INT64, cols.FLOAT, cols.DOUBLE])
def test_groupby_partitioned_path_random(ctx, kt, at):
    rng = np.random.default_rng(900 + 31 * kt + at)
    for n, hi, hint in ((1, 100, 50), (5000, 100, 100), (300_000, 1 << 40, 300_000), (300_000, 40_000, 2_000_000)):
        for null_prob in (0.0, 0.3):
            top = hi if kt == cols.DOUBLE else min(hi, int(np.iinfo(cols.NP_OF[kt]).max))
            key = rand_column(rng, kt, n, 0.05, lo=0, hi=top, special=(kt == cols.DOUBLE))
            arg = rand_column(rng, at, n, null_prob, special=True)
            args = [arg] * len(HP_AGGS)
            a = run_groupby("gpu", [key], HP_AGGS, args, ctx, expected_groups=hint, options=HP)
            b = run_groupby(ORACLE, [key], HP_AGGS, args)
            assert_group_equal(a, b, HP_AGGS, rtol=RTOL, scales=abs_sum_scales([key], HP_AGGS, args))


def test_groupby_partitioned_path_is_taken(ctx):
    """The options do route the call through the partition kernels (launch count of the context goes up by the
    partition / aggregate / import launches), and int64 sums wrap exactly like the reference's."""
    from arrow1_b200 import capi
    rng = np.random.default_rng(5)
    n = 200_000
    key = Column.from_numpy(rng.integers(0, 1 << 62, size=n, dtype=np.int64) % 50_000)
    arg = Column.from_numpy(rng.integers(-2**63, 2**63 - 1, size=n, dtype=np.int64), rng.random(n) > 0.1)
    aggs = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_MEAN), AggSpec(cols.HASH_COUNT)]
    capi.lib().b200_ctx_set_kernel_timing(ctx._h, 1)
    a = run_groupby("gpu", [key], aggs, [arg] * 3, ctx, expected_groups=50_000, options=HP)
    import ctypes as C
    ms, cnt = C.c_double(0), C.c_int64(0)
    capi.lib().b200_ctx_kernel_time(ctx._h, b"groupby_partition", C.byref(ms), C.byref(cnt))
    capi.lib().b200_ctx_set_kernel_timing(ctx._h, 0)
    assert cnt.value >= 2, "partition kernels did not run"
    b = run_groupby(ORACLE, [key], aggs, [arg] * 3)
    assert_group_equal(a, b, aggs, rtol=RTOL)


def test_groupby_partitioned_path_multi_key_and_no_aggs(ctx):
    rng = np.random.default_rng(12)
    n = 120_000
    k0 = rand_column(rng, cols.UINT8, n, 0.1, lo=0, hi=200)
    k1 = rand_column(rng, cols.INT16, n, 0.1, lo=-300, hi=300)
    k2 = rand_column(rng, cols.BOOL, n, 0.1)
    arg = rand_column(rng, cols.DOUBLE, n, 0.2)
    aggs = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_COUNT)]
    a = run_groupby("gpu", [k0, k1, k2], aggs, [arg] * 2, ctx, expected_groups=100_000, options=HP)
    b = run_groupby(ORACLE, [k0, k1, k2], aggs, [arg] * 2)
    assert_group_equal(a, b, aggs, rtol=RTOL, scales=abs_sum_scales([k0, k1, k2], aggs, [arg] * 2))
    a = run_groupby("gpu", [k0, k1], [], [], ctx, expected_groups=100_000, options=HP)
    b = run_groupby(ORACLE, [k0, k1], [], [])
    assert_group_equal(a, b, [], rtol=RTOL)


def test_groupby_partitioned_path_special_keys_and_skew(ctx):
    """Null / all-ones 64-bit keys (special slots) are deferred to the generic kernel; a key that holds half of the
    rows makes one partition far larger than the others (cut into several work items); a hint far below the real
    cardinality overflows the shared tables (rows deferred) and the image list."""
    rng = np.random.default_rng(13)
    n = 400_000
    v = rng.integers(0, 1 << 50, size=n, dtype=np.int64)
    v[rng.random(n) < 0.5] = 42
    v[rng.random(n) < 0.01] = -1
    key = Column.from_numpy(v, rng.random(n) > 0.02)
    arg = Column.from_numpy(rng.random(n) * 100.0, rng.random(n) > 0.05)
    aggs = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_MEAN), AggSpec(cols.HASH_COUNT, 1)]
    b = run_groupby(ORACLE, [key], aggs, [arg] * 3)
    for hint in (n, 1000):
        a = run_groupby("gpu", [key], aggs, [arg] * 3, ctx, expected_groups=hint, options=HP)
        assert_group_equal(a, b, aggs, rtol=RTOL, scales=abs_sum_scales([key], aggs, [arg] * 3))


def test_groupby_partitioned_path_streaming_selected_and_rows(ctx):
    """Several batches into one handle (the import folds into groups of earlier batches), a selection bitmap, and
    explicit row numbers (ascending: taken; shuffled: falls back to the generic kernels) — all equal one batch."""
    from arrow1_b200 import capi
    rng = np.random.default_rng(14)
    n = 300_000
    kv = rng.integers(0, 60_000, size=n).astype(np.int64) * 1_000_003
    key = Column.from_numpy(kv)
    arg = Column.from_numpy(rng.random(n), rng.random(n) > 0.1)
    aggs = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_COUNT)]
    want = run_groupby(ORACLE, [key], aggs, [arg] * 2)
    scales = abs_sum_scales([key], aggs, [arg] * 2)

    def finish(gb):
        out = gb.finalize()
        gb.close()
        res_aggs = [[(c.to_host().values().copy(), c.to_host().valid()) for c in a] for a in out.aggs]
        res_keys = [(c.to_host().values().copy(), c.to_host().valid()) for c in out.keys]
        from helpers import GroupRes
        return GroupRes(out.num_groups, res_aggs, res_keys)

    def handle(hint):
        gb = capi.GroupBy(ctx, [key.type], aggs, [arg.type] * 2, hint)
        for k, v_ in HP.items():
            gb.set_option(k, v_)
        return gb

    # three batches
    gb = handle(60_000)
    cuts = [0, 100_000, 100_001, n]
    dkey, darg = key.to_device(), arg.to_device()
    for lo, hi in zip(cuts[:-1], cuts[1:]):
        gb.consume([Column(dkey.type, dkey.data, None, hi - lo, lo, 0)],
                   [Column(darg.type, darg.data, darg.validity, hi - lo, lo, -1)] * 2, row_base=lo)
    assert_group_equal(finish(gb), want, aggs, rtol=RTOL, scales=scales)
    # explicit row numbers
    for shuffled in (False, True):
        gb = handle(60_000)
        rows = np.arange(n, dtype=np.uint64) * 3 + 7
        if shuffled:
            perm = rng.permutation(n)
            k2, a2, rows = Column.from_numpy(kv[perm]), Column.from_numpy(arg.values()[perm], arg.valid()[perm]), rows[perm]
        else:
            k2, a2 = key, arg
        gb.consume_rows([k2.to_device()], [a2.to_device()] * 2, torch.from_numpy(rows.view(np.int64)).cuda())
        assert_group_equal(finish(gb), want, aggs, rtol=RTOL, scales=scales)
    # selection
    sel = rng.random(n) < 0.6
    kept = np.flatnonzero(sel)
    want_sel = run_groupby(ORACLE, [Column.from_numpy(kv[kept])], aggs, [Column.from_numpy(arg.values()[kept], arg.valid()[kept])] * 2)
    gb = handle(60_000)
    gb.consume_selected([key.to_device()], [arg.to_device()] * 2, Column.from_numpy(sel).to_device())
    assert_group_equal(finish(gb), want_sel, aggs, rtol=RTOL)


@pytest.mark.parametrize("hp", [False, True])
def test_groupby_first_rows_before_and_after_finalize(ctx, hp):
    """b200_groupby_first_rows right after finalize reuses finalize's ranking; after further batches (or with no
    finalize at all) it ranks the groups itself.  Either way: the first row of every group, ascending."""
    from arrow1_b200 import capi
    rng = np.random.default_rng(41)
    n = 200_000
    kv = rng.integers(0, 30_000, size=n).astype(np.int64) * 7919
    arg = Column.from_numpy(rng.random(n))
    aggs = [AggSpec(cols.HASH_SUM)]

    def expect(upto, base):
        _, first = np.unique(kv[:upto], return_index=True)
        return np.sort(first).astype(np.uint64) + np.uint64(base)

    gb = capi.GroupBy(ctx, [cols.INT64], aggs, [cols.DOUBLE], 30_000)
    if hp:
        for k, v_ in HP.items():
            gb.set_option(k, v_)
    dkey, darg = Column.from_numpy(kv).to_device(), arg.to_device()
    half = n // 2
    gb.consume([Column(dkey.type, dkey.data, None, half, 0, 0)], [Column(darg.type, darg.data, None, half, 0, 0)], row_base=1000)
    np.testing.assert_array_equal(gb.first_rows().cpu().numpy().view(np.uint64), expect(half, 1000))   # no finalize yet
    out = gb.finalize()
    np.testing.assert_array_equal(gb.first_rows().cpu().numpy().view(np.uint64), expect(half, 1000))   # finalize's ranking
    np.testing.assert_array_equal(out.keys[0].to_host().values(), kv[(expect(half, 1000) - np.uint64(1000)).astype(np.int64)])
    gb.consume([Column(dkey.type, dkey.data, None, n - half, half, 0)], [Column(darg.type, darg.data, None, n - half, half, 0)],
               row_base=1000 + half)
    np.testing.assert_array_equal(gb.first_rows().cpu().numpy().view(np.uint64), expect(n, 1000))      # cache dropped
    out = gb.finalize()
    np.testing.assert_array_equal(gb.first_rows().cpu().numpy().view(np.uint64), expect(n, 1000))
    np.testing.assert_array_equal(out.keys[0].to_host().values(), kv[(expect(n, 1000) - np.uint64(1000)).astype(np.int64)])
    gb.close()


# ---- keys wider than 64 packed bits / more than 4 key columns (FixedWidthKeyEncoder takes any, hash_aggregate.cc:142-196)
@pytest.mark.parametrize("kts", [(cols.INT64, cols.INT64), (cols.INT64, cols.INT32, cols.UINT8),
                                 (cols.DOUBLE, cols.BOOL, cols.INT16, cols.UINT64, cols.FLOAT),
                                 (cols.UINT8,) * 6, (cols.INT32,) * 3])
def test_groupby_wide_keys(ctx, kts):
    rng = np.random.default_rng(sum(kts) + len(kts))
    for n, hi in ((1, 3), (3000, 3), (200_000, 40)):
        keys = [rand_column(rng, t, n, 0.1 if i % 2 == 0 else 0.0, lo=0, hi=(1 if t == cols.BOOL else hi),
                            special=(t in (cols.FLOAT, cols.DOUBLE))) for i, t in enumerate(kts)]
        arg = rand_column(rng, cols.DOUBLE, n, 0.2)
        iarg = rand_column(rng, cols.INT32, n, 0.1)
        aggs = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_COUNT), AggSpec(cols.HASH_MIN_MAX), AggSpec(cols.HASH_MEAN), AggSpec(cols.HASH_SUM)]
        args = [arg, arg, arg, arg, iarg]
        a = run_groupby("gpu", keys, aggs, args, ctx, want_ids=True)
        b = run_groupby(ORACLE, keys, aggs, args, want_ids=True)
        assert_group_equal(a, b, aggs, rtol=RTOL, scales=abs_sum_scales(keys, aggs, args))


def test_groupby_wide_keys_streaming_and_selection(ctx):
    """Two 64-bit keys over several batches (the key table grows and keeps the ids of earlier batches) and through a
    selection; Grouper ids continue across calls."""
    from arrow1_b200 import capi
    from helpers import GroupRes
    rng = np.random.default_rng(99)
    n = 150_000
    k0 = Column.from_numpy(rng.integers(0, 300, size=n).astype(np.int64) * (1 << 40))
    k1 = Column.from_numpy(rng.integers(-50, 50, size=n).astype(np.int64), rng.random(n) > 0.05)
    arg = Column.from_numpy(rng.random(n), rng.random(n) > 0.1)
    aggs = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_COUNT)]
    want = run_groupby(ORACLE, [k0, k1], aggs, [arg] * 2, want_ids=True)
    d0, d1, da = k0.to_device(), k1.to_device(), arg.to_device()
    gb = capi.GroupBy(ctx, [k0.type, k1.type], aggs, [arg.type] * 2)
    gr = capi.GroupBy(ctx, [k0.type, k1.type], [], [])
    ids = []
    for lo, hi in ((0, 1000), (1000, 70_000), (70_000, n)):
        sl = lambda c, nc: Column(c.type, c.data, c.validity, hi - lo, lo, nc)   # noqa: E731
        gb.consume([sl(d0, 0), sl(d1, -1)], [sl(da, -1)] * 2, row_base=lo)
        ids.append(gr.grouper_consume([sl(d0, 0), sl(d1, -1)]).cpu().numpy().view(np.uint32).copy())
    out = gb.finalize()
    gb.close()
    gr.close()
    a = GroupRes(out.num_groups, [[(c.to_host().values().copy(), c.to_host().valid()) for c in ag] for ag in out.aggs],
                 [(c.to_host().values().copy(), c.to_host().valid()) for c in out.keys], np.concatenate(ids))
    assert_group_equal(a, GroupRes(want.num_groups, want.aggs, want.keys, want.ids), aggs, rtol=RTOL)
    sel = rng.random(n) < 0.5
    kept = np.flatnonzero(sel)
    f = lambda c: Column.from_numpy(c.values()[kept].copy(), c.valid()[kept].copy() if c.validity is not None else None)   # noqa: E731
    want_sel = run_groupby(ORACLE, [f(k0), f(k1)], aggs, [f(arg)] * 2)
    got = capi.group_by(ctx, [d0, d1], aggs, [da] * 2, selection=Column.from_numpy(sel).to_device())
    a = GroupRes(got.num_groups, [[(c.to_host().values().copy(), c.to_host().valid()) for c in ag] for ag in got.aggs],
                 [(c.to_host().values().copy(), c.to_host().valid()) for c in got.keys])
    assert_group_equal(a, want_sel, aggs, rtol=RTOL)


# ---- other consumers of the Grouper: MakeGroupings / ApplyGroupings, unique, value_counts
def test_make_and_apply_groupings_match_reference(ctx):
    from arrow1_b200 import capi, compute
    if ORACLE != "reference":
        pytest.skip("needs oracle/_ref")
    # the documented example of api_aggregate.h:340-383
    ids = np.array([2, 2, 5, 5, 2, 3], dtype=np.uint32)
    off, idx = compute.MakeGroupings(torch.from_numpy(ids.view(np.int32)).cuda(), 8, ctx)
    assert off.cpu().tolist() == [0, 0, 0, 3, 4, 4, 6, 6, 6] and idx.cpu().tolist() == [0, 1, 4, 5, 2, 3]
    rng = np.random.default_rng(8)
    for n, g in ((0, 3), (1, 1), (1000, 7), (300_000, 5000), (300_000, 3)):
        ids = rng.integers(0, max(g - 1, 1), size=n).astype(np.uint32)   # the last group stays empty
        want_off, want_idx = oracle.make_groupings(ids, g)
        dids = torch.from_numpy(ids.view(np.int32)).cuda()
        off, idx = compute.MakeGroupings(dids, g, ctx)
        np.testing.assert_array_equal(off.cpu().numpy(), want_off)
        np.testing.assert_array_equal(idx.cpu().numpy(), want_idx)
        if n:
            vals = Column.from_numpy(rng.random(n), rng.random(n) > 0.1)
            _, taken = compute.ApplyGroupings((off, idx), vals.to_device(), ctx)
            h = taken.to_host()
            np.testing.assert_array_equal(h.valid(), vals.valid()[want_idx])
            np.testing.assert_array_equal(h.values()[h.valid()], vals.values()[want_idx][h.valid()])
    with pytest.raises(capi.B200Error) as e:
        compute.MakeGroupings(torch.tensor([0, 9, 1], dtype=torch.int32).cuda(), 4, ctx)
    assert e.value.code == 7


@pytest.mark.parametrize("t", [cols.BOOL, cols.INT8, cols.UINT16, cols.INT32, cols.INT64, cols.FLOAT, cols.DOUBLE])
def test_unique_and_value_counts_match_reference(ctx, t):
    from arrow1_b200 import compute
    if ORACLE != "reference":
        pytest.skip("needs oracle/_ref")
    rng = np.random.default_rng(60 + t)
    for n, nulls in ((1, 0.0), (5000, 0.0), (200_000, 0.1)):
        col = rand_column(rng, t, n, nulls, lo=0, hi=(1 if t == cols.BOOL else 90))
        (wv, wvalid), wc = oracle.value_counts(col, True)
        uv, _ = oracle.value_counts(col, False)
        got_u = compute.Unique(col.to_device(), ctx).to_host()
        np.testing.assert_array_equal(got_u.valid(), uv[1])
        np.testing.assert_array_equal(got_u.values()[uv[1]].view(np.uint8), uv[0][uv[1]].view(np.uint8))
        keys, counts = compute.ValueCounts(col.to_device(), ctx)
        hk = keys.to_host()
        np.testing.assert_array_equal(hk.valid(), wvalid)
        np.testing.assert_array_equal(hk.values()[wvalid].view(np.uint8), wv[wvalid].view(np.uint8))
        np.testing.assert_array_equal(counts.cpu().numpy(), wc)


@pytest.mark.parametrize("encode_nulls", [False, True])
@pytest.mark.parametrize("t", [cols.BOOL, cols.UINT8, cols.INT32, cols.INT64, cols.DOUBLE])
def test_dictionary_encode_matches_reference(ctx, t, encode_nulls):
    """"dictionary_encode" (vector_hash.cc:165-226): indices, their validity and the dictionary, both
    DictionaryEncodeOptions::NullEncodingBehavior modes; the null first met in the middle (ids behind it shift)."""
    from arrow1_b200 import compute
    if ORACLE != "reference":
        pytest.skip("needs oracle/_ref")
    rng = np.random.default_rng(160 + t)
    for n, nulls in ((1, 0.0), (33, 0.2), (5000, 0.0), (200_000, 0.05)):
        col = rand_column(rng, t, n, nulls, lo=0, hi=(1 if t == cols.BOOL else 200))
        widx, wvalid, (dv, dvalid) = oracle.dictionary_encode(col, encode_nulls)
        idx, dictionary = compute.DictionaryEncode(col.to_device(), encode_nulls, ctx)
        hi, hd = idx.to_host(), dictionary.to_host()
        np.testing.assert_array_equal(hi.valid(), wvalid)
        np.testing.assert_array_equal(hi.values()[wvalid], widx[wvalid])
        if not encode_nulls:
            assert (hi.values()[~wvalid] == 0).all()
        np.testing.assert_array_equal(hd.valid(), dvalid)
        np.testing.assert_array_equal(hd.values()[dvalid].view(np.uint8), dv[dvalid].view(np.uint8))


# ---- variable-length (binary / utf8) group keys (VarLengthKeyEncoder, hash_aggregate.cc:236-311)
def _random_strings(rng, n, n_distinct, null_p, max_len=14):
    vocab = [b"", b"a", b"a\x00", b"a\x00b", b"ab", b"abc", b"abcdefgh", b"abcdefghi", b"abcdefgh\x00"]
    while len(vocab) < n_distinct:
        ln = int(rng.integers(0, max_len + 1))
        vocab.append(bytes(rng.integers(0, 256, size=ln, dtype=np.uint8)))
    vocab = vocab[:n_distinct]
    pick = rng.integers(0, len(vocab), size=n)
    nulls = rng.random(n) < null_p
    return [None if nulls[i] else vocab[pick[i]] for i in range(n)]


@pytest.mark.parametrize("shape", ["s", "s+i32", "i64+s+s", "s,no nulls"])
def test_groupby_string_keys_match_reference(ctx, shape):
    if ORACLE != "reference":
        pytest.skip("needs oracle/_ref")
    rng = np.random.default_rng(300 + len(shape))
    n = 60_000
    s1 = Column.from_strings(_random_strings(rng, n, 700, 0.0 if "no nulls" in shape else 0.05))
    keys = [s1]
    if shape == "s+i32":
        keys = [s1, rand_column(rng, cols.INT32, n, 0.05, lo=0, hi=5)]
    if shape == "i64+s+s":
        keys = [rand_column(rng, cols.INT64, n, 0.02, lo=0, hi=3), s1, Column.from_strings(_random_strings(rng, n, 9, 0.3, max_len=40))]
    arg = rand_column(rng, cols.DOUBLE, n, 0.1)
    aggs = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_COUNT), AggSpec(cols.HASH_MIN_MAX)]
    args = [arg] * 3
    got = run_groupby("gpu", keys, aggs, args, ctx, want_ids=True)
    want = run_groupby(ORACLE, keys, aggs, args, want_ids=True)
    assert_group_equal(got, want, aggs, rtol=RTOL, scales=abs_sum_scales(keys, aggs, args, ORACLE))
    np.testing.assert_array_equal(got.ids, want.ids)


def test_groupby_string_keys_streaming_and_edge_cases(ctx):
    """Batches into one handle (later batches bring new strings: the key heap and the id table grow), a sliced
    column (non-zero offset), an all-empty-string column, an empty batch."""
    if ORACLE != "reference":
        pytest.skip("needs oracle/_ref")
    from arrow1_b200 import capi
    from helpers import GroupRes, _host_key
    rng = np.random.default_rng(77)
    n = 90_000
    items = _random_strings(rng, n, 40, 0.1)[: n // 3] + _random_strings(rng, n, 5000, 0.1)[n // 3:]
    key = Column.from_strings(items)
    arg = Column.from_numpy(rng.random(n), rng.random(n) > 0.1)
    aggs = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_COUNT)]
    want = run_groupby(ORACLE, [key], aggs, [arg] * 2)
    gb = capi.GroupBy(ctx, [cols.BINARY], aggs, [arg.type] * 2, 0)
    dkey, darg = key.to_device(), arg.to_device()
    cuts = [0, 10, 10, n // 3 + 17, n]
    for lo, hi in zip(cuts[:-1], cuts[1:]):
        gb.consume([Column(dkey.type, dkey.data, dkey.validity, hi - lo, lo, -1, dkey.extra)],
                   [Column(darg.type, darg.data, darg.validity, hi - lo, lo, -1)] * 2, row_base=lo)
    out = gb.finalize()
    gb.close()
    got = GroupRes(out.num_groups, [[(c.to_host().values().copy(), c.to_host().valid()) for c in a] for a in out.aggs],
                   [_host_key(c) for c in out.keys])
    assert_group_equal(got, want, aggs, rtol=RTOL, scales=abs_sum_scales([key], aggs, [arg] * 2, ORACLE))
    # all rows the empty string / all null / no rows
    for items in ([""] * 100, [None] * 100, []):
        k = Column.from_strings(items)
        a = Column.from_numpy(np.arange(len(items), dtype=np.int64))
        cnt = [AggSpec(cols.HASH_COUNT)]
        got = run_groupby("gpu", [k], cnt, [a], ctx)
        if items:
            assert_group_equal(got, run_groupby(ORACLE, [k], cnt, [a]), cnt)
        else:
            assert got.num_groups == 0


def test_string_hash_kernels_and_host_groupby(ctx):
    """unique / value_counts / dictionary_encode on a binary column against the reference, and compute.GroupBy from
    HOST columns keyed by strings (the value bytes travel with the offsets)."""
    from arrow1_b200 import compute
    from helpers import GroupRes
    if ORACLE != "reference":
        pytest.skip("needs oracle/_ref")
    rng = np.random.default_rng(19)
    n = 40_000
    col = Column.from_strings(_random_strings(rng, n, 300, 0.08))
    (wv, wvalid), wc = oracle.value_counts(col, True)
    (uv, uvalid), _ = oracle.value_counts(col, False)
    got_u = compute.Unique(col.to_device(), ctx).to_host()
    assert got_u.strings() == [x if v else None for x, v in zip(uv, uvalid)]
    keys, counts = compute.ValueCounts(col.to_device(), ctx)
    assert keys.to_host().strings() == [x if v else None for x, v in zip(wv, wvalid)]
    np.testing.assert_array_equal(counts.cpu().numpy(), wc)
    for encode_nulls in (False, True):
        widx, wivalid, (dv, dvalid) = oracle.dictionary_encode(col, encode_nulls)
        idx, dictionary = compute.DictionaryEncode(col.to_device(), encode_nulls, ctx)
        hi = idx.to_host()
        np.testing.assert_array_equal(hi.valid(), wivalid)
        np.testing.assert_array_equal(hi.values()[wivalid], widx[wivalid])
        assert dictionary.to_host().strings() == [x if v else None for x, v in zip(dv, dvalid)]
    arg = Column.from_numpy(rng.random(n))
    aggs = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_COUNT)]
    res = compute.GroupBy([arg, arg], [col], aggs, ctx)
    want = run_groupby(ORACLE, [col], aggs, [arg, arg])
    assert_group_equal(GroupRes(res.num_groups, res.aggs, res.keys), want, aggs, rtol=RTOL)
