"""Take parity: the CUDA gather must equal the reference's PrimitiveTake bit for bit —
data (zero under nulls), the always-present validity bitmap, the exact null count — and raise its IndexError."""
import numpy as np
import pytest

import oracle
from arrow1_b200 import capi, columns as cols
from arrow1_b200.capi import B200Error
from arrow1_b200.columns import Column
from gen import ALL_NUMERIC, rand_column

pytestmark = pytest.mark.gpu
ORACLE = "reference" if oracle.have_reference() else "port"
INDEX_TYPES = [(cols.INT8, np.int8), (cols.UINT8, np.uint8), (cols.INT16, np.int16), (cols.UINT16, np.uint16),
               (cols.INT32, np.int32), (cols.UINT32, np.uint32), (cols.INT64, np.int64), (cols.UINT64, np.uint64)]


def check(ctx, values: Column, indices: Column, kind=ORACLE):
    want = oracle.take(values, indices, True, kind=kind)
    got = capi.take(ctx, values.to_device(), indices.to_device(), True).to_host()
    n = indices.length
    assert got.length == want.length == n
    assert got.null_count == want.null_count
    np.testing.assert_array_equal(got.valid(), want.valid())
    bits = cols.BITS[values.type]
    nbytes = (n + 7) // 8 if bits == 1 else n * (bits // 8)
    if bits == 1:
        np.testing.assert_array_equal(got.values(), want.values())
    else:
        assert got.data[:nbytes].tobytes() == want.data[:nbytes].tobytes()   # incl. zero bytes under nulls


@pytest.mark.parametrize("t", ALL_NUMERIC + [cols.BOOL])
def test_take_random_values_every_type(ctx, t):
    # vector_selection_test.cc TakeTest: random values x random indices, null_prob {0, .1, .5}
    rng = np.random.default_rng(900 + t)
    for nv, ni in ((1, 5), (100, 1000), (4099, 70001)):
        for vnull, inull in ((0.0, 0.0), (0.1, 0.0), (0.0, 0.3), (0.5, 0.5)):
            if t == cols.BOOL:
                values = Column.from_numpy(rng.random(nv + 3) < 0.5, (rng.random(nv + 3) >= vnull) if vnull else None, offset=3)
            else:
                values = rand_column(rng, t, nv, vnull, offset=3, special=True)
            ix = rng.integers(0, nv, size=ni + 7)
            indices = Column.from_numpy(ix.astype(np.int32), (rng.random(ni + 7) >= inull) if inull else None, offset=7)
            check(ctx, values, indices)


@pytest.mark.parametrize("it,dt", INDEX_TYPES)
def test_take_every_index_type(ctx, it, dt):
    rng = np.random.default_rng(50 + it)
    nv = min(100, int(np.iinfo(dt).max))
    values = rand_column(rng, cols.INT64, nv, 0.2)
    ix = rng.integers(0, nv, size=5000).astype(dt)
    check(ctx, values, Column.from_numpy(ix, rng.random(5000) > 0.1))
    check(ctx, values, Column.from_numpy(ix))


def test_take_golden_vectors(ctx):
    # vector_selection_test.cc: TakeNumeric cases ([7,8,9] x [0,1,0]; [null,8,9] x [0,1,0]; [7,8,9] x [null,1,0];
    # [null,8,9] x [null,1,0]) and empty indices
    v = np.array([7, 8, 9], dtype=np.int8)
    for vvalid, ivalid, want_v, want_valid in (
            (None, None, [7, 8, 7], [1, 1, 1]),
            ([0, 1, 1], None, [0, 8, 0], [0, 1, 0]),
            (None, [0, 1, 1], [0, 8, 7], [0, 1, 1]),
            ([0, 1, 1], [0, 1, 1], [0, 8, 0], [0, 1, 0])):
        values = Column.from_numpy(v, None if vvalid is None else np.array(vvalid, bool))
        indices = Column.from_numpy(np.array([0, 1, 0], dtype=np.int32), None if ivalid is None else np.array(ivalid, bool))
        got = capi.take(ctx, values.to_device(), indices.to_device(), True).to_host()
        np.testing.assert_array_equal(got.values(), np.array(want_v, dtype=np.int8))
        np.testing.assert_array_equal(got.valid(), np.array(want_valid, bool))
        assert got.null_count == 3 - sum(want_valid)
        check(ctx, values, indices, "port")
    empty = capi.take(ctx, Column.from_numpy(v).to_device(), Column.from_numpy(np.zeros(0, dtype=np.int32)).to_device(), True)
    assert empty.length == 0 and empty.null_count == 0


def test_take_out_of_bounds(ctx):
    # vector_selection_test.cc: indices [0, 9, 0] into 3 values -> IndexError; negative index -> IndexError;
    # a NULL out-of-range index is fine
    values = Column.from_numpy(np.array([7, 8, 9], dtype=np.int64)).to_device()
    for bad, dt in (([0, 9, 0], np.int8), ([0, -1, 0], np.int64), ([3], np.uint64)):
        with pytest.raises(B200Error) as e:
            capi.take(ctx, values, Column.from_numpy(np.array(bad, dtype=dt)).to_device(), True)
        assert e.value.code == 7 and ("Index %d out of bounds" % [b for b in bad if b < 0 or b >= 3][0]) in str(e.value)
        with pytest.raises(oracle.OracleError) as oe:
            oracle.take(values.to_host(), Column.from_numpy(np.array(bad, dtype=dt)), True, kind=ORACLE)
        assert oe.value.code == 7
    ok = capi.take(ctx, values, Column.from_numpy(np.array([0, 9, 1], dtype=np.int32), np.array([1, 0, 1], bool)).to_device(), True).to_host()
    np.testing.assert_array_equal(ok.valid(), [True, False, True])
    with pytest.raises(B200Error) as e:
        capi.take(ctx, values, Column.from_numpy(np.array([0.0])).to_device(), True)
    assert e.value.code == 3


def test_take_applies_sort_indices(ctx):
    """SortIndices -> Take materialises the sorted column (nulls last), the use the survey names."""
    rng = np.random.default_rng(4)
    col = rand_column(rng, cols.DOUBLE, 200_000, 0.05, special=True)
    dev = col.to_device()
    idx = capi.sort_indices(ctx, dev, 0)
    got = capi.take(ctx, dev, idx, True).to_host()
    ref_idx = oracle.sort_indices(col, 0, 0, kind=ORACLE)
    want = oracle.take(col, Column.from_numpy(ref_idx), True, kind=ORACLE)
    assert got.null_count == want.null_count == col.length - int(col.valid().sum())
    np.testing.assert_array_equal(got.valid(), want.valid())
    assert got.data[: col.length * 8].tobytes() == want.data[: col.length * 8].tobytes()


def _chunked(rng, whole: Column, cuts, offsets=True):
    """Slices of a host column as separate Columns (their own buffers, odd offsets, some without a validity)."""
    vals, valid = whole.values(), whole.valid()
    out = []
    for a, b in zip(cuts[:-1], cuts[1:]):
        off = int(rng.integers(0, 13)) if offsets else 0
        v = np.concatenate([np.ones(off, dtype=vals.dtype), vals[a:b]])
        m = np.concatenate([np.ones(off, dtype=np.bool_), valid[a:b]])
        out.append(Column.from_numpy(v, None if m[off:].all() and (a + b) % 2 else m, offset=off))
    return out


@pytest.mark.parametrize("t", [cols.INT8, cols.UINT16, cols.FLOAT, cols.INT64, cols.DOUBLE, cols.BOOL])
def test_take_chunked_values_equal_take_of_the_concatenation(ctx, t):
    """b200_take_chunked == TakeCA (vector_selection.cc:1921-1944): Concatenate(chunks) then array_take — the
    oracle runs on the whole column, the device gathers from the chunks where they lie.  Empty chunks (leading,
    middle, trailing), a chunk without validity, odd offsets, null indices."""
    rng = np.random.default_rng(4000 + t)
    for nv, cuts in ((10, [0, 0, 3, 3, 10, 10]), (5000, [0, 1, 2, 4097, 4097, 5000]), (70_000, None)):
        if cuts is None:
            cuts = [0] + sorted(rng.integers(0, nv, size=37).tolist()) + [nv]
        if t == cols.BOOL:
            whole = Column.from_numpy(rng.random(nv) < 0.5, rng.random(nv) >= 0.2)
        else:
            whole = rand_column(rng, t, nv, 0.2, special=True)
        chunks = [c.to_device() for c in _chunked(rng, whole, cuts)]
        ni = 3 * nv + 5
        indices = Column.from_numpy(rng.integers(0, nv, size=ni).astype(np.int64), rng.random(ni) >= 0.1)
        want = oracle.take(whole, indices, True, kind=ORACLE)
        got = capi.take(ctx, chunks, indices.to_device(), True).to_host()
        assert got.length == want.length and got.null_count == want.null_count
        np.testing.assert_array_equal(got.valid(), want.valid())
        np.testing.assert_array_equal(got.values(), want.values())   # zeros under nulls on both sides
        # a single "chunk" and the plain call agree
        one = capi.take(ctx, [whole.to_device()], indices.to_device(), True).to_host()
        np.testing.assert_array_equal(one.values(), want.values())
        np.testing.assert_array_equal(one.valid(), want.valid())


def test_take_chunked_many_chunks_and_bounds(ctx):
    rng = np.random.default_rng(77)
    nv = 200_000
    whole = rand_column(rng, cols.INT32, nv, 0.05)
    cuts = [0] + sorted(rng.integers(0, nv, size=2999).tolist()) + [nv]      # 3000 chunks: boundaries in global memory
    chunks = [c.to_device() for c in _chunked(rng, whole, cuts, offsets=False)]
    ix = rng.integers(0, nv, size=300_000).astype(np.uint32)
    want = oracle.take(whole, Column.from_numpy(ix), True, kind=ORACLE)
    got = capi.take(ctx, chunks, Column.from_numpy(ix).to_device(), True).to_host()
    np.testing.assert_array_equal(got.values(), want.values())
    np.testing.assert_array_equal(got.valid(), want.valid())
    assert got.null_count == want.null_count
    # the bound is the TOTAL length
    bad = Column.from_numpy(np.array([0, nv - 1, nv, 1], dtype=np.int64)).to_device()
    with pytest.raises(B200Error) as e:
        capi.take(ctx, chunks, bad, True)
    assert e.value.code == 7 and "Index %d out of bounds" % nv in str(e.value)
    ok = capi.take(ctx, chunks, bad, False).to_host()     # NoBoundsCheck: the row comes out null
    np.testing.assert_array_equal(ok.valid()[[0, 1, 3]], whole.valid()[[0, nv - 1, 1]])
    assert not ok.valid()[2]
    # all chunks empty: every index is out of bounds
    empty = [Column.from_numpy(np.zeros(0, dtype=np.int32)).to_device() for _ in range(3)]
    with pytest.raises(B200Error):
        capi.take(ctx, empty, Column.from_numpy(np.array([0], dtype=np.int8)).to_device(), True)
    # chunks of different types
    with pytest.raises(B200Error) as e:
        capi.take(ctx, [chunks[0], Column.from_numpy(np.zeros(4, dtype=np.int64)).to_device()], bad, False)
    assert e.value.code == 3


def test_compute_take_accepts_host_chunks(ctx):
    from arrow1_b200 import compute
    rng = np.random.default_rng(5)
    whole = rand_column(rng, cols.DOUBLE, 30_000, 0.1, special=True)
    chunks = _chunked(rng, whole, [0, 10_000, 10_001, 30_000])
    indices = Column.from_numpy(rng.integers(0, 30_000, size=50_000).astype(np.int32))
    want = oracle.take(whole, indices, True, kind=ORACLE)
    got = compute.Take(chunks, indices, ctx=ctx)
    assert not got.on_device and got.null_count == want.null_count
    np.testing.assert_array_equal(got.valid(), want.valid())
    assert got.values().tobytes() == want.values().tobytes()
