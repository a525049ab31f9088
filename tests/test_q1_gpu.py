"""Config C5: the TPC-H Q1-shaped pipeline (mask producer -> Filter of every used column -> GroupBy on
(returnflag, linestatus)) through the C ABI vs the same pipeline through the oracle (the reference's
Filter and GroupBy; the mask is the reference's less_equal semantics, scalar_compare.cc: C++ `<=` on the
physical values, stated here with numpy).  Bit-exact group order / keys / counts, float sums and means
within 1e-12.  Also the mask producer on its own for every numeric type and operator."""
import numpy as np
import pytest
import torch

import oracle
from arrow1_b200 import capi, columns as cols, compute, synth
from arrow1_b200.columns import AggSpec, Column
from gen import rand_column
from helpers import GroupRes, assert_group_equal

pytestmark = pytest.mark.gpu
ORACLE = "reference" if oracle.have_reference() else "port"
NP_CMP = {"equal": np.equal, "not_equal": np.not_equal, "greater": np.greater, "greater_equal": np.greater_equal,
          "less": np.less, "less_equal": np.less_equal}


@pytest.mark.parametrize("t", [cols.INT8, cols.UINT8, cols.INT16, cols.UINT16, cols.INT32, cols.UINT32, cols.INT64,
                               cols.UINT64, cols.FLOAT, cols.DOUBLE])
def test_compare_scalar_matches_cxx_operators(ctx, t):
    rng = np.random.default_rng(40 + t)
    for n in (1, 31, 32, 127, 128, 129, 100_003):
        col = rand_column(rng, t, n, 0.0, lo=0, hi=50, special=(t in (cols.FLOAT, cols.DOUBLE)))
        v = col.values()
        scalar = v[n // 2] if n > 1 else v[0]
        if v.dtype.kind == "f" and np.isnan(scalar):
            scalar = v.dtype.type(3.0)
        for op, fn in NP_CMP.items():
            got = capi.compare_scalar(ctx, col.to_device(), op, scalar)
            bits = np.unpackbits(got.data.cpu().numpy()[: (n + 7) // 8], bitorder="little")[:n].astype(bool)
            with np.errstate(invalid="ignore"):
                np.testing.assert_array_equal(bits, fn(v, scalar), err_msg="%s type %d n %d" % (op, t, n))


def test_compare_scalar_sliced_input(ctx):
    rng = np.random.default_rng(8)
    v = rng.integers(-100, 100, size=5000).astype(np.int32)
    col = Column.from_numpy(v, offset=37)
    got = capi.compare_scalar(ctx, col.to_device(), "less", 5)
    n = col.length
    bits = np.unpackbits(got.data.cpu().numpy()[: (n + 7) // 8], bitorder="little")[:n].astype(bool)
    np.testing.assert_array_equal(bits, v[37:] < 5)


def _host(c):
    h = c.to_host()
    return h.values().copy(), h.valid()


@pytest.mark.parametrize("n", [1000, 250_000])
def test_q1_pipeline_matches_oracle(ctx, n):
    host = synth.Gen(None).q1_lineitem(n, start=12345)
    dev = synth.Gen("cuda:0").q1_lineitem(n, start=12345)
    for h, d in zip(host, dev):   # both back ends generate identical bytes
        nb = h.length * cols.BITS[h.type] // 8
        assert h.data[:nb].tobytes() == d.data[:nb].cpu().numpy().tobytes()
    ship, rf, ls, qty, price, disc = dev
    got = compute.Q1Pipeline(ship, rf, ls, [qty, price, disc], ctx)

    hship, hrf, hls, hqty, hprice, hdisc = host
    mask = Column.from_numpy(hship.values() <= compute.Q1_SHIPDATE_CUTOFF)
    filt = [oracle.filter(c, mask, 0, kind=ORACLE) for c in (hrf, hls, hqty, hprice, hdisc)]
    fcols = [Column.from_numpy(f.values().copy()) for f in filt]
    fcols.append(Column.from_numpy(fcols[3].values() * (1.0 - fcols[4].values())))   # disc_price, IEEE double ops
    specs = compute.q1_aggregates()
    want = oracle.groupby([fcols[0], fcols[1]], [a for _, a in specs], [fcols[2 + i] for i, _ in specs], kind=ORACLE)

    a = GroupRes(got.num_groups, [[_host(c) for c in ag] for ag in got.aggs], [_host(k) for k in got.keys])
    b = GroupRes(want.num_groups, want.aggs, want.keys)
    assert a.num_groups == b.num_groups and a.num_groups <= 4   # A/F, N/F, N/O, R/F
    assert_group_equal(a, b, [s for _, s in specs], rtol=1e-12)
    # the Filter folded into the aggregation (b200_groupby_consume_selected): same groups, same order, same values
    got2 = compute.Q1PipelineSelected(ship, rf, ls, [qty, price, disc], ctx)
    a2 = GroupRes(got2.num_groups, [[_host(c) for c in ag] for ag in got2.aggs], [_host(k) for k in got2.keys])
    assert_group_equal(a2, b, [s for _, s in specs], rtol=1e-12)
    # everything in one pass over the inputs (b200_groupby_consume_fused; below 8192 rows through the separate operations)
    got3 = compute.Q1PipelineFused(ship, rf, ls, [qty, price, disc], ctx)
    a3 = GroupRes(got3.num_groups, [[_host(c) for c in ag] for ag in got3.aggs], [_host(k) for k in got3.keys])
    assert_group_equal(a3, b, [s for _, s in specs], rtol=1e-12)
    # counts, keys and group order of the fused and the unfused form are the same bits
    for x, y in zip(a3.keys, a.keys):
        np.testing.assert_array_equal(x[0], y[0])
    np.testing.assert_array_equal(a3.aggs[6][0][0], a.aggs[6][0][0])


@pytest.mark.parametrize("shape", ["fast", "nine_groups", "int_keys_nulls", "no_predicate"])
def test_groupby_consume_fused_equals_separate_operations(ctx, shape):
    """b200_groupby_consume_fused == compare_scalar -> Filter(DROP) of every column -> arithmetic -> GroupBy, for the
    one-pass kernel's shape, for a 9th group (the kernel gives up, the call is redone unfused), for other key types and
    nullable arguments (evaluated through the separate operations), and without a predicate."""
    rng = np.random.default_rng(hash(shape) % 1000)
    n = 300_000
    groups = 9 if shape == "nine_groups" else 6
    if shape == "int_keys_nulls":
        k0 = Column.from_numpy(rng.integers(0, 50, size=n).astype(np.int32), rng.random(n) > 0.05)
        k1 = Column.from_numpy(rng.integers(0, 3, size=n).astype(np.uint8))
        hint = 200
    else:
        k0 = Column.from_numpy(rng.integers(0, groups, size=n).astype(np.uint8))
        k1 = Column.from_numpy(np.zeros(n, dtype=np.uint8) + 7)
        hint = 8
    nulls = shape == "int_keys_nulls"
    x = Column.from_numpy(rng.random(n) * 1e3, (rng.random(n) > 0.1) if nulls else None)
    y = Column.from_numpy(rng.random(n), (rng.random(n) > 0.1) if nulls else None)
    date = Column.from_numpy(rng.integers(0, 1000, size=n).astype(np.int32))
    pred = None if shape == "no_predicate" else (date, "greater", 300)
    aggs = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_MEAN), AggSpec(cols.HASH_COUNT), AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_MEAN)]
    gb = capi.GroupBy(ctx, [k0.type, k1.type], aggs, [cols.DOUBLE] * 5, hint)
    dk0, dk1, dx, dy, dd = (c.to_device() for c in (k0, k1, x, y, date))
    derived = ("add", dx, "multiply", 2.5, dy)
    gb.consume_fused([dk0, dk1], [dx, dy, dx, derived, derived], None if pred is None else (dd, pred[1], pred[2]))
    got = gb.finalize()
    gb.close()
    keep = np.ones(n, dtype=bool) if pred is None else (date.values() > 300)
    idx = np.flatnonzero(keep)

    def filt(c):
        return Column.from_numpy(c.values()[idx].copy(), c.valid()[idx].copy() if c.validity is not None else None)

    fx, fy = filt(x), filt(y)
    dv = fx.values() + 2.5 * fy.values()
    fd = Column.from_numpy(dv, (fx.valid() & fy.valid()) if nulls else None)
    want = oracle.groupby([filt(k0), filt(k1)], aggs, [fx, fy, fx, fd, fd], kind=ORACLE)
    a = GroupRes(got.num_groups, [[_host(c) for c in ag] for ag in got.aggs], [_host(k) for k in got.keys])
    b = GroupRes(want.num_groups, want.aggs, want.keys)
    assert_group_equal(a, b, aggs, rtol=1e-12)


NP_ARITH = {"add": np.add, "subtract": np.subtract, "multiply": np.multiply}


@pytest.mark.parametrize("t", [cols.INT8, cols.UINT8, cols.INT16, cols.UINT16, cols.INT32, cols.UINT32, cols.INT64,
                               cols.UINT64, cols.FLOAT, cols.DOUBLE])
def test_arithmetic_matches_wrapping_and_ieee_semantics(ctx, t):
    """add / subtract / multiply (scalar_arithmetic.cc:69-190, unchecked): integers wrap, floats are single IEEE
    operations; validity = AND of the operands'; array-array, array-scalar and scalar-array forms; sliced inputs."""
    rng = np.random.default_rng(500 + t)
    for n in (1, 33, 1024, 100_003):
        a = rand_column(rng, t, n, 0.1, offset=3, special=(t in (cols.FLOAT, cols.DOUBLE)))
        b = rand_column(rng, t, n, 0.0 if n == 33 else 0.2, offset=5, special=(t in (cols.FLOAT, cols.DOUBLE)))
        va, vb = a.values(), b.values()
        s = va[n // 2]
        if va.dtype.kind == "f" and not np.isfinite(s):
            s = va.dtype.type(1.5)
        da, db = a.to_device(), b.to_device()
        for op, fn in NP_ARITH.items():
            with np.errstate(all="ignore"):
                for got, want, valid in (
                        (capi.arithmetic(ctx, op, da, db), fn(va, vb), a.valid() & b.valid()),
                        (capi.arithmetic(ctx, op, da, s), fn(va, s), a.valid()),
                        (capi.arithmetic(ctx, op, s, db), fn(s, vb), b.valid())):
                    h = got.to_host()
                    want = want.astype(va.dtype)
                    gv = h.values()
                    if va.dtype.kind == "f":   # a NaN result carries the GPU's canonical payload, not the x86 one
                        nan = np.isnan(want)
                        np.testing.assert_array_equal(np.isnan(gv), nan)
                        gv, want = gv[~nan], want[~nan]
                    assert gv.tobytes() == want.tobytes(), "%s type %d n %d" % (op, t, n)
                    np.testing.assert_array_equal(h.valid(), valid)


def test_arithmetic_errors(ctx):
    from arrow1_b200.capi import B200Error
    a = Column.from_numpy(np.arange(10, dtype=np.int32)).to_device()
    b = Column.from_numpy(np.arange(11, dtype=np.int32)).to_device()
    c = Column.from_numpy(np.arange(10, dtype=np.int64)).to_device()
    with pytest.raises(B200Error) as e:
        capi.arithmetic(ctx, "add", a, b)
    assert e.value.code == 4 and "same length" in str(e.value)
    with pytest.raises(B200Error) as e:
        capi.arithmetic(ctx, "add", a, c)
    assert e.value.code == 10


@pytest.mark.parametrize("kt,groups", [(cols.UINT8, 5), (cols.INT32, 700), (cols.INT64, 200_000)])
def test_groupby_consume_selected_equals_groupby_of_filtered_columns(ctx, kt, groups):
    """b200_groupby_consume_selected == GroupBy over Filter(column, selection) of every column, for the tiny, the
    shared-table and the global-table paths (min_max and nulls included)."""
    rng = np.random.default_rng(groups)
    n = 400_000
    key = Column.from_numpy(rng.integers(0, groups, size=n).astype(cols.NP_OF[kt]))
    arg = Column.from_numpy(rng.random(n) * 1e6 - 5e5, rng.random(n) > 0.03)
    sel_bits = rng.random(n) < 0.6
    sel = Column.from_numpy(sel_bits)
    for aggs in ([AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_COUNT), AggSpec(cols.HASH_MEAN)],
                 [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_MIN_MAX), AggSpec(cols.HASH_COUNT, 1)]):
        got = capi.group_by(ctx, [key.to_device()], aggs, [arg.to_device()] * len(aggs), expected_groups=groups,
                            selection=sel.to_device())
        fk = oracle.filter(key, sel, 0, kind=ORACLE)
        fa = oracle.filter(arg, sel, 0, kind=ORACLE)
        fkc = Column.from_numpy(fk.values().copy())
        fac = Column.from_numpy(fa.values().copy(), fa.valid().copy())
        want = oracle.groupby([fkc], aggs, [fac] * len(aggs), kind=ORACLE)
        a = GroupRes(got.num_groups, [[_host(c) for c in ag] for ag in got.aggs], [_host(k) for k in got.keys])
        b = GroupRes(want.num_groups, want.aggs, want.keys)
        assert_group_equal(a, b, aggs, rtol=1e-12)
