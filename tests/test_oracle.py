"""Pins oracle/oracle.cc (the restated algorithm) against the UNMODIFIED reference built from
/root/reference (oracle/_ref) on randomised inputs shaped like the reference's own property
tests.  CPU only."""
import numpy as np
import pytest

import oracle
from arrow1_b200 import columns as cols
from arrow1_b200.columns import AggSpec
from gen import ALL_NUMERIC, rand_column, rand_mask
from helpers import multi_key_cases, assert_filter_equal, assert_group_equal, run_filter, run_groupby, run_sort

needs_ref = pytest.mark.skipif(not oracle.have_reference(), reason="oracle/_ref not built")


@needs_ref
@pytest.mark.parametrize("t", [cols.BOOL] + ALL_NUMERIC)
def test_filter_port_matches_reference(t):
    rng = np.random.default_rng(1234 + t)
    # vector_selection_test.cc:1618-1636: null_prob x true_prob matrix, length 1024, mis-aligned slices
    for null_prob in (0.0, 0.01, 0.1, 0.999, 1.0):
        for true_prob in (0.0, 0.1, 0.999, 1.0):
            for (voff, moff) in ((0, 0), (3, 4)):
                for mask_null in (0.0, 0.05):
                    values = rand_column(rng, t, 1024, null_prob, offset=voff)
                    mask = rand_mask(rng, 1024, true_prob, mask_null, offset=moff)
                    for null_sel in (0, 1):
                        a = run_filter("port", values, mask, null_sel)
                        b = run_filter("reference", values, mask, null_sel)
                        assert_filter_equal(a, b, bit_exact=True)


@needs_ref
def test_filter_unknown_null_count_port_matches_reference():
    # vector_selection_test.cc:445-455: null_count unknown (-1) on both sides
    rng = np.random.default_rng(7)
    values = rand_column(rng, cols.INT32, 500, 0.2)
    mask = rand_mask(rng, 500, 0.5, 0.1)
    values.null_count = -1
    mask.null_count = -1
    for null_sel in (0, 1):
        assert_filter_equal(run_filter("port", values, mask, null_sel), run_filter("reference", values, mask, null_sel))


@needs_ref
@pytest.mark.parametrize("t", ALL_NUMERIC)
def test_sort_port_matches_reference(t):
    rng = np.random.default_rng(99 + t)
    for n in (0, 1, 100, 5000):
        for null_prob in (0.0, 0.1, 0.5, 1.0):
            for kw in (dict(), dict(lo=-100 if np.dtype(cols.NP_OF[t]).kind in "if" else 0, hi=100), dict(special=True)):
                col = rand_column(rng, t, n, null_prob, offset=3, **kw)
                for order in (0, 1):
                    a = run_sort("port", col, order)
                    b = run_sort("reference", col, order)
                    np.testing.assert_array_equal(a, b)


@needs_ref
@pytest.mark.parametrize("kt", [cols.BOOL, cols.UINT8, cols.INT32, cols.INT64, cols.FLOAT, cols.DOUBLE])
@pytest.mark.parametrize("at", [cols.INT8, cols.UINT16, cols.INT32, cols.UINT64, cols.INT64, cols.FLOAT, cols.DOUBLE])
def test_groupby_port_matches_reference(kt, at):
    rng = np.random.default_rng(5 + 31 * kt + at)
    for n in (0, 1, 1 << 10, 1 << 12):
        for null_prob in (0.0, 0.01, 0.5, 1.0):
            key = rand_column(rng, kt, n, 0.05, lo=0, hi=100, special=(kt in (cols.FLOAT, cols.DOUBLE)))
            arg = rand_column(rng, at, n, null_prob, special=True)
            aggs = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_COUNT), AggSpec(cols.HASH_COUNT, 1),
                    AggSpec(cols.HASH_MIN_MAX), AggSpec(cols.HASH_MIN_MAX, 1), AggSpec(cols.HASH_MEAN)]
            args = [arg] * len(aggs)
            a = run_groupby("port", [key], aggs, args, want_ids=True)
            b = run_groupby("reference", [key], aggs, args, want_ids=True)
            # both are sequential in row order: even float sums must be bit-identical
            assert_group_equal(a, b, aggs, rtol=0)


@needs_ref
def test_groupby_multi_key_port_matches_reference():
    rng = np.random.default_rng(11)
    n = 4096
    k0 = rand_column(rng, cols.UINT8, n, 0.1, lo=0, hi=3)
    k1 = rand_column(rng, cols.INT16, n, 0.1, lo=-2, hi=2)
    k2 = rand_column(rng, cols.BOOL, n, 0.1)
    arg = rand_column(rng, cols.DOUBLE, n, 0.2)
    aggs = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_MIN_MAX), AggSpec(cols.HASH_COUNT)]
    a = run_groupby("port", [k0, k1, k2], aggs, [arg] * 3, want_ids=True)
    b = run_groupby("reference", [k0, k1, k2], aggs, [arg] * 3, want_ids=True)
    assert_group_equal(a, b, aggs, rtol=0)


def test_oracle_builds_and_exports():
    oracle.build()
    lib, _ = oracle._lib("port")
    for name in ("oracle_filter", "oracle_filter_count", "oracle_sort_indices", "oracle_groupby", "oracle_version"):
        assert hasattr(lib, name)


@needs_ref
@pytest.mark.parametrize("t", ALL_NUMERIC + [cols.BOOL])
def test_take_port_matches_reference(t):
    """oracle_take (the restatement of PrimitiveTake, vector_selection.cc:273-526) against compute::Take itself."""
    from arrow1_b200.columns import Column
    rng = np.random.default_rng(1200 + t)
    for nv, ni in ((1, 4), (50, 500), (1000, 5000)):
        for vnull, inull in ((0.0, 0.0), (0.2, 0.0), (0.0, 0.3), (0.4, 0.4)):
            if t == cols.BOOL:
                values = Column.from_numpy(rng.random(nv + 2) < 0.5, (rng.random(nv + 2) >= vnull) if vnull else None, offset=2)
            else:
                values = rand_column(rng, t, nv, vnull, offset=2, special=True)
            for dt in (np.int8, np.uint16, np.int32, np.uint64):
                hi = min(nv, int(np.iinfo(dt).max))
                ix = rng.integers(0, hi, size=ni + 5).astype(dt)
                indices = Column.from_numpy(ix, (rng.random(ni + 5) >= inull) if inull else None, offset=5)
                a = oracle.take(values, indices, True, kind="port")
                b = oracle.take(values, indices, True, kind="reference")
                assert a.length == b.length and a.null_count == b.null_count
                assert a.data_bytes() == b.data_bytes() and a.validity_bytes() == b.validity_bytes()
    values = rand_column(rng, t, 20, 0.0) if t != cols.BOOL else Column.from_numpy(rng.random(20) < 0.5)
    for bad in ([0, 1, 50], [-1]):
        for kind in ("port", "reference"):
            with pytest.raises(oracle.OracleError) as e:
                oracle.take(values, Column.from_numpy(np.array(bad, dtype=np.int64)), True, kind=kind)
            assert e.value.code == 7


@needs_ref
def test_sort_multi_key_port_matches_reference_record_batch_and_table():
    # vector_sort_test.cc TestTableSortIndices / TestRecordBatchSortIndices: nulls and NaNs last per key, stable
    rng = np.random.default_rng(31)
    for n in (0, 1, 7, 3000):
        for keys, orders in multi_key_cases(rng, n):
            port = oracle.sort_indices_multi(keys, orders)
            np.testing.assert_array_equal(port, oracle.sort_indices_multi(keys, orders, kind="reference"))
            np.testing.assert_array_equal(port, oracle.sort_indices_multi(keys, orders, kind="reference", as_table=True, n_chunks=3))
    # nine keys: past kMaxRadixSortKeys the RecordBatch form switches to MultipleKeyRecordBatchSorter
    keys = [rand_column(rng, cols.INT8, 2000, 0.1, lo=0, hi=2) for _ in range(9)]
    orders = [i % 2 for i in range(9)]
    np.testing.assert_array_equal(oracle.sort_indices_multi(keys, orders), oracle.sort_indices_multi(keys, orders, kind="reference"))
    # one key == SortIndices of the column
    np.testing.assert_array_equal(oracle.sort_indices_multi(keys[:1], [1]), oracle.sort_indices(keys[0], 1))


def test_sort_multi_key_golden_vectors():
    from helpers import MULTI_KEY_GOLDEN
    for keys, orders, want in MULTI_KEY_GOLDEN():
        np.testing.assert_array_equal(oracle.sort_indices_multi(keys, orders), np.array(want, dtype=np.uint64))
