"""SortIndices parity: the CUDA radix sort must produce the reference's exact stable permutation."""
import numpy as np
import pytest

import oracle
from arrow1_b200 import capi, columns as cols
from arrow1_b200.capi import B200Error
from arrow1_b200.columns import Column
from conftest import rand_u64
from gen import ALL_NUMERIC, rand_column
from helpers import run_sort

pytestmark = pytest.mark.gpu
ORACLE = "reference" if oracle.have_reference() else "port"


@pytest.mark.parametrize("t", ALL_NUMERIC)
def test_sort_random(ctx, t):
    # vector_sort_test.cc:550-624 (length 100, null_prob {0,.1,.5,1}, both orders) + larger sizes
    rng = np.random.default_rng(777 + t)
    lo = -100 if np.dtype(cols.NP_OF[t]).kind in "if" else 0
    for n in (1, 100, 4095, 4097, 70001):
        for null_prob in (0.0, 0.1, 0.5, 1.0):
            for kw in (dict(), dict(lo=lo, hi=100), dict(special=True)):
                col = rand_column(rng, t, n, null_prob, offset=5, **kw)
                for order in (0, 1):
                    a = run_sort("gpu", col, order, 0, ctx)
                    b = run_sort(ORACLE, col, order)
                    np.testing.assert_array_equal(a, b, err_msg="type %d n %d null %.1f %s order %d" % (t, n, null_prob, kw, order))


def test_sort_int64_extremes_with_nulls(ctx):
    """INT64_MAX / INT64_MIN rows share the radix key of nulls: the 257th bin must split them."""
    rng = np.random.default_rng(5)
    n = 50000
    info = np.iinfo(np.int64)
    for t, dt in ((cols.INT64, np.int64), (cols.UINT64, np.uint64)):
        ii = np.iinfo(dt)
        v = rng.choice(np.array([ii.min, ii.max, 0, 1, ii.max - 1], dtype=dt), size=n)
        valid = rng.random(n) > 0.3
        col = Column.from_numpy(v, valid)
        for order in (0, 1):
            np.testing.assert_array_equal(run_sort("gpu", col, order, 0, ctx), run_sort(ORACLE, col, order))


def test_sort_index_base(ctx):
    rng = np.random.default_rng(6)
    col = rand_column(rng, cols.DOUBLE, 1000, 0.1)
    a = run_sort("gpu", col, 0, 1 << 40, ctx)
    b = run_sort(ORACLE, col, 0, 1 << 40)
    np.testing.assert_array_equal(a, b)


def test_sort_boolean_not_implemented(ctx):
    from arrow1_b200 import capi
    col = Column.from_numpy(np.array([True, False])).to_device()
    with pytest.raises(capi.B200Error) as ei:
        capi.sort_indices(ctx, col, 0)
    assert ei.value.code == 10


def _is_sorted_stable(values, valid, idx, order):
    """SortComparator as a spec (vector_sort_test.cc:102-125): total order with row-index tie break."""
    v = values[idx]
    ok = valid[idx]
    n_valid = int(valid.sum())
    assert ok[:n_valid].all() and not ok[n_valid:].any()
    isnan = np.isnan(v[:n_valid]) if v.dtype.kind == "f" else np.zeros(n_valid, dtype=bool)
    n_num = n_valid - int(isnan.sum())
    assert not isnan[:n_num].any()
    a, b = v[:n_num][:-1], v[:n_num][1:]
    ia, ib = idx[:n_num][:-1], idx[:n_num][1:]
    if order == 0:
        assert ((a < b) | ((a == b) & (ia < ib))).all()
    else:
        assert ((a > b) | ((a == b) & (ia < ib))).all()
    for seg in (idx[n_num:n_valid], idx[n_valid:]):
        assert (np.diff(seg.astype(np.int64)) > 0).all()


@pytest.mark.parametrize("kind", ["int64", "float64"])
def test_sort_config2_properties_large(ctx, kind):
    """BASELINE config 2 shape (reduced to 20 M rows for test time; bench.py runs 100 M):
    sortedness + stability + null/NaN placement checked as properties, and the first 2 M rows
    prefix-independent spot check against the oracle on a 2 M-row column."""
    n = 20_000_000
    with np.errstate(over="ignore"):
        if kind == "int64":
            values = rand_u64(4, n).view(np.int64)
        else:
            values = (rand_u64(5, n) >> np.uint64(11)).astype(np.float64) * (2.0 ** -53) * 1e14
        valid = (rand_u64(6, n) % np.uint64(100)) != 0
    col = Column.from_numpy(values, valid)
    idx = run_sort("gpu", col, 0, 0, ctx)
    assert len(np.unique(idx)) == n  # a permutation
    _is_sorted_stable(values, valid, idx, 0)
    m = 2_000_000
    small = Column.from_numpy(values[:m], valid[:m])
    np.testing.assert_array_equal(run_sort("gpu", small, 0, 0, ctx), run_sort(ORACLE, small, 0))


def test_sort_pairs_stable(ctx):
    from arrow1_b200 import capi
    import torch
    rng = np.random.default_rng(8)
    n = 300000
    keys = rng.integers(0, 1000, size=n, dtype=np.uint64)
    payload = rng.integers(0, 1 << 62, size=n, dtype=np.uint64)
    k, p = capi.sort_pairs(ctx, torch.from_numpy(keys.view(np.int64)).cuda(), torch.from_numpy(payload.view(np.int64)).cuda(),
                           want_keys=True)
    order = np.argsort(keys, kind="stable")
    np.testing.assert_array_equal(k.cpu().numpy().view(np.uint64), keys[order])
    np.testing.assert_array_equal(p.cpu().numpy().view(np.uint64), payload[order])


def test_sort_keys_transform_is_order_preserving(ctx):
    from arrow1_b200 import capi
    rng = np.random.default_rng(9)
    col = rand_column(rng, cols.DOUBLE, 10000, 0.1, special=True)
    keys = capi.sort_keys(ctx, col.to_device(), 0).cpu().numpy().view(np.uint64)
    idx = run_sort(ORACLE, col, 0)
    assert (np.diff(keys[idx].astype(np.float64)) >= 0).all() or (keys[idx][:-1] <= keys[idx][1:]).all()


@pytest.mark.parametrize("n_parts", [1, 2, 3, 8, 16])
def test_range_partition_is_stable_and_matches_numpy_image(ctx, n_parts):
    """b200_sort_range_partition (the multi-GPU sample sort's local step): dest = number of splitters <=
    radix key; every destination's (key, row) pairs contiguous and in row order."""
    from arrow1_b200 import capi
    from test_dist_cpu import radix_keys
    rng = np.random.default_rng(70 + n_parts)
    for n in (1, 33, 4095, 4096, 4097, 100_003):
        f = rng.integers(-50, 50, size=n).astype(np.float64)
        f[rng.integers(0, n, max(n // 50, 1))] = np.nan
        for col in (Column.from_numpy(f, rng.random(n) > 0.05),
                    Column.from_numpy(rng.integers(-2**63, 2**63 - 1, size=n, dtype=np.int64), rng.random(n) > 0.05)):
            for order in (0, 1):
                k, cls = radix_keys(col, order)
                spl = np.sort(rng.choice(k, size=n_parts - 1)) if n_parts > 1 else np.zeros(0, dtype=np.uint64)
                spl = np.minimum(spl, np.uint64(2**64 - 2))
                keys, rows, counts = capi.sort_range_partition(ctx, col.to_device(), order, 1000, [int(x) for x in spl], n_parts)
                dest = np.searchsorted(spl, k, side="right")
                p = np.argsort(dest, kind="stable")
                want_rows = (np.arange(n, dtype=np.uint64) + np.uint64(1000)) | (cls.astype(np.uint64) << np.uint64(62))
                assert counts == [int((dest == q).sum()) for q in range(n_parts)]
                np.testing.assert_array_equal(keys.cpu().numpy().view(np.uint64), k[p])
                np.testing.assert_array_equal(rows.cpu().numpy().view(np.uint64), want_rows[p])


def test_sort_indices_chunked_equals_sort_of_concatenation(ctx):
    """ChunkedArraySorter semantics (vector_sort.cc:657-832): global stable order over the logical
    concatenation, NaNs then nulls of every chunk last; genuine maxima stay ahead of them."""
    from arrow1_b200 import capi
    rng = np.random.default_rng(321)
    n = 50_000
    f = rng.integers(-20, 20, size=n).astype(np.float64)
    f[rng.integers(0, n, 900)] = np.nan
    f[rng.integers(0, n, 200)] = np.inf
    u = rng.integers(0, 2**64 - 1, size=n, dtype=np.uint64)
    u[rng.integers(0, n, 300)] = np.uint64(2**64 - 1)
    u[rng.integers(0, n, 300)] = np.uint64(2**64 - 2)
    for vals in (f, u):
        valid = rng.random(n) > 0.05
        whole = Column.from_numpy(vals, valid)
        cuts = [0, 1, 17000, 17000, 42000, n]
        chunks = [Column.from_numpy(vals[a:b].copy(), valid[a:b].copy()).to_device() for a, b in zip(cuts[:-1], cuts[1:])]
        for order in (0, 1):
            got = capi.sort_indices_chunked(ctx, chunks, order).cpu().numpy().view(np.uint64)
            np.testing.assert_array_equal(got, run_sort(ORACLE, whole, order))


def test_range_partition_push_and_sort_pairs32_match_the_unfused_forms(ctx):
    """b200_sort_range_partition_count/_push into per-destination buffers at arbitrary offsets == the runs of
    b200_sort_range_partition; b200_sort_pairs32 == b200_sort_pairs with the payload widened."""
    import torch
    from arrow1_b200 import capi
    rng = np.random.default_rng(99)
    n = 123_457
    f = np.floor(rng.random(n) * 3000.0) - 1500.0
    f[rng.integers(0, n, 700)] = np.nan
    col = Column.from_numpy(f, rng.random(n) > 0.03).to_device()
    keys_all = capi.sort_keys(ctx, col, 0).cpu().numpy().view(np.uint64)
    spl = [int(x) for x in np.quantile(keys_all[keys_all < np.uint64(2**63)], [0.3, 0.8]).astype(np.uint64)]
    base = 1000
    ks, rs, counts = capi.sort_range_partition(ctx, col, 0, base, spl, 3)
    plan = capi.RangePartitionPush(ctx, col, 0, spl, 3)
    assert plan.counts == counts
    offs = [5, 0, 7]
    kb = [capi.PeerBuffer(ctx, (c + o) * 8 + 64) for c, o in zip(counts, offs)]
    rb = [capi.PeerBuffer(ctx, (c + o) * 4 + 64) for c, o in zip(counts, offs)]
    try:
        plan.push(base, [b.ptr for b in kb], [b.ptr for b in rb], offs, rows32=True)
        start = 0
        for q in range(3):
            k = kb[q].tensor(torch.int64, counts[q], offs[q] * 8)
            r = rb[q].tensor(torch.int32, counts[q], offs[q] * 4)
            want_k, want_r = ks[start:start + counts[q]], rs[start:start + counts[q]]
            assert torch.equal(k, want_k)
            want32 = (want_r & ((1 << 30) - 1)) | (((want_r >> 62) & 3) << 30)
            assert torch.equal(r.to(torch.int64) & 0xFFFFFFFF, want32)
            start += counts[q]
        # pairs with a 32-bit payload
        k = kb[2].tensor(torch.int64, counts[2], offs[2] * 8)
        r = rb[2].tensor(torch.int32, counts[2], offs[2] * 4)
        sk32, sp32 = capi.sort_pairs32(ctx, k, r, want_keys=True)
        sk, sp = capi.sort_pairs(ctx, k.clone(), (r.to(torch.int64) & 0xFFFFFFFF).contiguous(), want_keys=True)
        assert torch.equal(sk32, sk) and torch.equal(sp32, sp)
    finally:
        for b in kb + rb:
            b.close()


def test_sort_indices_multi_key_golden_vectors(ctx):
    from helpers import MULTI_KEY_GOLDEN
    for keys, orders, want in MULTI_KEY_GOLDEN():
        got = capi.sort_indices_multi(ctx, [k.to_device() for k in keys], orders).cpu().numpy().view(np.uint64)
        np.testing.assert_array_equal(got, np.array(want, dtype=np.uint64))


def test_sort_indices_multi_key_matches_oracle(ctx):
    """b200_sort_indices_multi == SortIndices(RecordBatch, sort_keys) of the reference: LSD over the keys against
    the reference's comparator sorts, every type, mixed orders, nulls / NaNs / -0.0 in every key, heavy ties."""
    from helpers import multi_key_cases
    rng = np.random.default_rng(8)
    for n in (1, 100, 70_001):
        for keys, orders in multi_key_cases(rng, n):
            want = oracle.sort_indices_multi(keys, orders, kind=ORACLE)
            got = capi.sort_indices_multi(ctx, [k.to_device() for k in keys], orders).cpu().numpy().view(np.uint64)
            np.testing.assert_array_equal(got, want)
    # every numeric type as the second key, offsets, wide value ranges
    for t in ALL_NUMERIC:
        first = rand_column(rng, cols.INT8, 50_000, 0.05, lo=0, hi=3)
        second = rand_column(rng, t, 50_000, 0.1, offset=5, special=True)
        for orders in ([0, 0], [1, 0], [0, 1]):
            want = oracle.sort_indices_multi([first, second], orders, kind=ORACLE)
            got = capi.sort_indices_multi(ctx, [first.to_device(), second.to_device()], orders).cpu().numpy().view(np.uint64)
            np.testing.assert_array_equal(got, want)
    # one key is the single-column sort; host columns go through compute.SortIndices
    from arrow1_b200 import compute
    one = rand_column(rng, cols.DOUBLE, 10_000, 0.1, special=True)
    np.testing.assert_array_equal(capi.sort_indices_multi(ctx, [one.to_device()], [1]).cpu().numpy().view(np.uint64),
                                  oracle.sort_indices(one, 1, kind=ORACLE))
    keys, orders = next(multi_key_cases(rng, 5000))
    np.testing.assert_array_equal(compute.SortIndices(keys, orders, ctx=ctx), oracle.sort_indices_multi(keys, orders, kind=ORACLE))


def test_sort_indices_multi_key_errors(ctx):
    a = Column.from_numpy(np.arange(10, dtype=np.int32)).to_device()
    b = Column.from_numpy(np.arange(11, dtype=np.int32)).to_device()
    with pytest.raises(B200Error) as e:
        capi.sort_indices_multi(ctx, [a, b], [0, 0])
    assert e.value.code == 4
    with pytest.raises(B200Error) as e:
        capi.sort_indices_multi(ctx, [], [])
    assert e.value.code == 4 and "one or more sort keys" in str(e.value)
    flag = Column.from_numpy(np.arange(10) % 2 == 0).to_device()
    with pytest.raises(B200Error) as e:
        capi.sort_indices_multi(ctx, [a, flag], [0, 0])
    assert e.value.code == 10


def test_sort_indices_multi_key_chunked_table(ctx):
    """b200_sort_indices_multi_chunked == SortIndices(Table) over multi-chunk columns (vector_sort_test.cc:886-980
    "Same data, several chunks"): the golden cases cut in two, and random tables cut at random places (empty
    chunks included) against the reference's MultipleKeyTableSorter on the whole columns."""
    from helpers import MULTI_KEY_GOLDEN, multi_key_cases

    def cut(col, cuts):
        vals, valid = col.values(), col.valid()
        return [Column.from_numpy(vals[a:b].copy(), None if (valid[a:b].all() and (a + b) % 2) else valid[a:b].copy()).to_device()
                for a, b in zip(cuts[:-1], cuts[1:])]
    for keys, orders, want in MULTI_KEY_GOLDEN():
        n = keys[0].length
        for cuts in ([0, n // 2, n], [0, 0, 1, n, n]):
            got = capi.sort_indices_multi(ctx, [cut(k, cuts) for k in keys], orders).cpu().numpy().view(np.uint64)
            np.testing.assert_array_equal(got, np.array(want, dtype=np.uint64))
    rng = np.random.default_rng(12)
    for n in (50, 40_000):
        for keys, orders in multi_key_cases(rng, n):
            cuts = [0] + sorted(rng.integers(0, n, size=5).tolist()) + [n]
            want = oracle.sort_indices_multi(keys, orders, kind=ORACLE, as_table=True, n_chunks=4) if ORACLE == "reference" \
                else oracle.sort_indices_multi(keys, orders)
            got = capi.sort_indices_multi(ctx, [cut(k, cuts) for k in keys], orders).cpu().numpy().view(np.uint64)
            np.testing.assert_array_equal(got, want)
    with pytest.raises(B200Error):
        a = Column.from_numpy(np.arange(10, dtype=np.int32)).to_device()
        capi.sort_indices_multi(ctx, [[a, a], [a]], [0, 0])


# ---- high word first (sort.cu HwState): 64-bit columns sorted by the upper key word, ties repaired afterwards -------
def _hw_cases(rng):
    n = 50_000
    rnd = rng.integers(-2**63, 2**63 - 1, size=n, dtype=np.int64)
    cases = {"random": rnd.copy()}
    pairs = rnd.copy()                       # many short runs: rows sharing the upper word, lower words in random order
    pairs[1::2] = (pairs[0::2] & ~np.int64(0xFFFFFFFF)) | rng.integers(0, 2**32, size=n // 2, dtype=np.int64)
    cases["pairs"] = rng.permutation(pairs)
    runs = rnd.copy()                        # runs of up to 32 rows (the repair's limit) and exact duplicates
    for start in range(0, 2000, 40):
        ln = int(rng.integers(2, 33))
        runs[start:start + ln] = (runs[start] & ~np.int64(0xFFFFFFFF)) | rng.integers(0, 4, size=ln, dtype=np.int64)
    cases["runs"] = rng.permutation(runs)
    long_run = rnd.copy()                    # one run of 33+ rows: the repair gives up, round 2 runs the ordinary plan
    long_run[:500] = (long_run[0] & ~np.int64(0xFFFFFFFF)) | rng.integers(0, 2**32, size=500, dtype=np.int64)
    cases["long run"] = rng.permutation(long_run)
    top = rnd.copy()                         # genuine keys in the run of the all-ones upper word, next to the nulls
    top[:20] = np.int64(-1) << np.int64(32) | rng.integers(0, 2**32, size=20, dtype=np.int64)
    top[20:25] = -1
    top[25:30] = np.iinfo(np.int64).max
    cases["top run"] = rng.permutation(top)
    return cases


@pytest.mark.parametrize("mode", [1, 0, -1])
@pytest.mark.parametrize("t", [cols.INT64, cols.UINT64, cols.DOUBLE])
def test_sort_indices_high_word_first(ctx, t, mode):
    """Forced (1), decided from the histograms (0) and off (-1): always the reference's permutation."""
    rng = np.random.default_rng(4242 + t)
    ctx.set_option("sort_high_word_first", mode)
    try:
        for name, v in _hw_cases(rng).items():
            vals = v.view(cols.NP_OF[t]).copy()
            if t == cols.DOUBLE:
                vals[:50] = np.nan
                vals[50:60] = -0.0
                vals[60:70] = 0.0
            for nulls in (0.0, 0.02):
                col = Column.from_numpy(vals, rng.random(len(vals)) >= nulls if nulls else None)
                for order in (0, 1):
                    want = oracle.sort_indices(col, order, kind=ORACLE)
                    got = capi.sort_indices(ctx, col.to_device(), order, index_base=7).cpu().numpy().view(np.uint64)
                    np.testing.assert_array_equal(got, want + np.uint64(7), err_msg="%s nulls=%s order=%d" % (name, nulls, order))
        ctx.synchronize()
    finally:
        ctx.set_option("sort_high_word_first", 0)


@pytest.mark.parametrize("mode", [1, 0, -1])
def test_sort_pairs32_high_word_first(ctx, mode):
    """(uint64 key, uint32 payload) pairs — the second key of a multi-key sort, the received pairs of the multi-GPU sort:
    stable by key whatever the plan; a long run of all-ones keys (nulls) keeps the ordinary plan."""
    import torch
    rng = np.random.default_rng(99)
    ctx.set_option("sort_high_word_first", mode)
    try:
        cases = _hw_cases(rng)
        nulls = cases["random"].copy()
        nulls[rng.random(len(nulls)) < 0.02] = -1          # 2 % all-ones keys
        cases["all-ones run"] = nulls
        for name, v in cases.items():
            keys = v.view(np.uint64)
            payload = rng.integers(0, 2**32, size=len(keys), dtype=np.uint32)
            want = payload[np.argsort(keys, kind="stable")].astype(np.uint64)
            _, got = capi.sort_pairs32(ctx, torch.from_numpy(keys.view(np.int64)).cuda(), torch.from_numpy(payload.view(np.int32)).cuda())
            np.testing.assert_array_equal(got.cpu().numpy().view(np.uint64), want, err_msg=name)
        ctx.synchronize()
    finally:
        ctx.set_option("sort_high_word_first", 0)


def test_sort_indices_plan_is_chosen_from_the_histograms(ctx):
    """Left to the estimate: random int64 takes the upper-key-word-first plan, float64 data (few distinct exponent
    words) and a small column do not, a column with one long run of equal upper words tries and falls back."""
    rng = np.random.default_rng(31)
    n = 300_000
    rnd = rng.integers(-2**63, 2**63 - 1, size=n, dtype=np.int64)
    long_run = rnd.copy()
    long_run[:1000] = (long_run[0] & ~np.int64(0xFFFFFFFF)) | rng.integers(0, 2**32, size=1000, dtype=np.int64)
    cases = [("random int64", Column.from_numpy(rnd, rng.random(n) > 0.01), (True, False)),
             ("float64 with few distinct upper words", Column.from_numpy(rng.integers(0, 1000, size=n).astype(np.float64) + rng.random(n) * 1e-9,
                                                                        rng.random(n) > 0.01), (False, False)),
             ("long run", Column.from_numpy(rng.permutation(long_run)), (True, True)),
             ("small", Column.from_numpy(rnd[:50_000].copy()), (False, False))]
    for name, col, plan in cases:
        scratch = ctx.empty(capi.sort_scratch_bytes(ctx, col.length))
        got = capi.sort_indices(ctx, col.to_device(), 0, scratch=scratch).cpu().numpy().view(np.uint64)
        np.testing.assert_array_equal(got, oracle.sort_indices(col, 0, kind=ORACLE), err_msg=name)
        assert capi.sort_indices_plan(ctx, scratch) == plan, name
