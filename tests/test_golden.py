"""Golden vectors transcribed from the reference's own tests (tests/golden/*.json), pushed through
the oracle port, the unmodified reference (when oracle/_ref is built) and — marked gpu — the CUDA
path through the C ABI."""
import numpy as np
import pytest

import oracle
from arrow1_b200 import columns as cols
from arrow1_b200.columns import AggSpec, Column
from conftest import TYPE_BY_NAME, column_from_json, expected_from_json, load_golden
from helpers import run_filter, run_groupby, run_sort

IMPLS = [
    pytest.param("port"),
    pytest.param("reference", marks=pytest.mark.skipif(not oracle.have_reference(), reason="oracle/_ref not built")),
    pytest.param("gpu", marks=pytest.mark.gpu),
]


def _ctx(impl, request):
    return request.getfixturevalue("ctx") if impl == "gpu" else None


def _coalesce_null_to_false(mask: Column) -> Column:
    v = mask.values() & mask.valid()
    return Column.from_numpy(v)


def _check_filter_case(impl, ctx, type_name, case):
    values = column_from_json(case["values"], type_name)
    mask = column_from_json(case["mask"], "bool")
    if "mask_slice" in case:
        off, ln = case["mask_slice"]
        parent = mask
        mask = Column(parent.type, parent.data, parent.validity, ln, off, -1 if parent.validity is not None else 0)
    exp_vals, exp_valid = expected_from_json(case["expected"], type_name)
    # EMIT_NULL against the literal expectation (vector_selection_test.cc:184-187)
    r = run_filter(impl, values, mask, 1, ctx)
    assert r.length == len(exp_vals)
    np.testing.assert_array_equal(r.valid, exp_valid)
    np.testing.assert_array_equal(r.values[exp_valid], exp_vals[exp_valid])
    # DROP == EMIT_NULL on the null-coalesced mask (vector_selection_test.cc:189-196)
    r_drop = run_filter(impl, values, mask, 0, ctx)
    r_ref = run_filter(impl, values, _coalesce_null_to_false(mask), 1, ctx)
    assert r_drop.length == r_ref.length
    np.testing.assert_array_equal(r_drop.valid, r_ref.valid)
    np.testing.assert_array_equal(r_drop.values[r_drop.valid], r_ref.values[r_ref.valid])


@pytest.mark.parametrize("impl", IMPLS)
def test_filter_numeric_golden(impl, request):
    g = load_golden("filter.json")
    ctx = _ctx(impl, request)
    for type_name in g["numeric_types"]:
        for case in g["numeric"]:
            _check_filter_case(impl, ctx, type_name, case)


@pytest.mark.parametrize("impl", IMPLS)
def test_filter_boolean_golden(impl, request):
    g = load_golden("filter.json")
    ctx = _ctx(impl, request)
    for case in g["boolean"]:
        _check_filter_case(impl, ctx, "bool", case)


@pytest.mark.parametrize("impl", IMPLS)
def test_filter_length_mismatch_is_invalid(impl, request):
    g = load_golden("filter.json")["length_mismatch"]
    ctx = _ctx(impl, request)
    values = column_from_json(g["values"], "int32")
    mask = column_from_json(g["mask"], "bool")
    for null_sel in (0, 1):
        with pytest.raises(Exception) as ei:
            run_filter(impl, values, mask, null_sel, ctx)
        assert getattr(ei.value, "code", None) == 4  # StatusCode::Invalid


@pytest.mark.parametrize("impl", IMPLS)
def test_sort_indices_golden(impl, request):
    g = load_golden("sort_indices.json")
    ctx = _ctx(impl, request)
    groups = [(g["real_types"], g["real"]), (g["integral_types"], g["integral"]), (["uint8"], g["uint8"]),
              (["int8"], g["int8"])]
    for types, cases in groups:
        for type_name in types:
            for case in cases:
                col = column_from_json(case["values"], type_name)
                out = run_sort(impl, col, case["order"], 0, ctx)
                np.testing.assert_array_equal(out, np.array(case["expected"], dtype=np.uint64),
                                              err_msg="%s %s" % (type_name, case))


def _batch(g):
    arg = column_from_json(g["batch"]["argument"], "double")
    key = column_from_json(g["batch"]["key"], "int64")
    return arg, key


def _cmp(col, expected, type_name):
    vals, valid = col
    ev, evalid = expected_from_json(expected, type_name)
    np.testing.assert_array_equal(valid, evalid)
    np.testing.assert_array_equal(vals[valid], ev[evalid])


@pytest.mark.parametrize("impl", IMPLS)
def test_group_by_golden(impl, request):
    g = load_golden("group_by.json")
    ctx = _ctx(impl, request)
    arg, key = _batch(g)
    r = run_groupby(impl, [key], [AggSpec(cols.HASH_SUM)], [arg], ctx)
    assert r.num_groups == 4
    _cmp(r.aggs[0][0], g["sum_only"]["hash_sum"], "double")
    _cmp(r.keys[0], g["sum_only"]["key_0"], "int64")

    r = run_groupby(impl, [key], [AggSpec(cols.HASH_MIN_MAX)], [arg], ctx)
    _cmp(r.aggs[0][0], g["min_max_only"]["min"], "double")
    _cmp(r.aggs[0][1], g["min_max_only"]["max"], "double")
    _cmp(r.keys[0], g["min_max_only"]["key_0"], "int64")

    # "passing an argument twice or also using it as a key is legal" (hash_aggregate_test.cc:575)
    aggs = [AggSpec(cols.HASH_COUNT), AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_SUM)]
    r = run_groupby(impl, [key], aggs, [arg, arg, key], ctx)
    e = g["count_and_sum"]
    _cmp(r.aggs[0][0], e["hash_count"], "int64")
    _cmp(r.aggs[1][0], e["hash_sum_f64"], "double")
    _cmp(r.aggs[2][0], e["hash_sum_key_i64"], "int64")
    _cmp(r.keys[0], e["key_0"], "int64")


def _grouper_sequence(impl, ctx, key_types, batches):
    """Successive Consume calls keep handing out ids in first-occurrence order."""
    if impl == "gpu":
        from arrow1_b200 import capi
        gb = capi.GroupBy(ctx, [TYPE_BY_NAME[t] for t in key_types], [], [])
        for b in batches:
            keys = [column_from_json(v, t).to_device() for v, t in zip(b["keys"], key_types)]
            ids = gb.grouper_consume(keys).cpu().numpy().view(np.uint32)
            np.testing.assert_array_equal(ids, np.array(b["ids"], dtype=np.uint32), err_msg=str((key_types, b)))
        gb.close()
        return
    # the CPU oracles are single-shot: replay the concatenation and compare the tail
    acc = [[] for _ in key_types]
    expected = []
    for b in batches:
        for i, v in enumerate(b["keys"]):
            acc[i] += v
        expected += b["ids"]
        keys = [column_from_json(v, t) for v, t in zip(acc, key_types)]
        r = oracle.groupby(keys, [], [], want_ids=True, kind=impl)
        np.testing.assert_array_equal(r.ids, np.array(expected, dtype=np.uint32), err_msg=str((key_types, b)))


@pytest.mark.parametrize("impl", IMPLS)
def test_grouper_golden(impl, request):
    g = load_golden("group_by.json")
    ctx = _ctx(impl, request)
    numeric = ["uint8", "int8", "uint16", "int16", "uint32", "int32", "uint64", "int64", "float", "double"]
    for seq in g["grouper_sequences"]:
        if seq["key_types"] == ["NUMERIC"]:
            for t in numeric:
                _grouper_sequence(impl, ctx, [t], seq["batches"])
        else:
            _grouper_sequence(impl, ctx, seq["key_types"], seq["batches"])


def _json_column(values, type_name):
    if type_name == "binary":
        return cols.Column.from_strings(values)
    return column_from_json(values, type_name)


@pytest.mark.parametrize("impl", IMPLS[1:])
def test_grouper_string_keys_golden(impl, request):
    """Grouper.StringKey / StringInt64Key / DoubleStringInt64Key of the reference's own tests (the C++ restatement in
    oracle/oracle.cc has no variable-length keys: the compiled reference and the CUDA path are checked)."""
    g = load_golden("group_by.json")
    ctx = _ctx(impl, request)
    for seq in g["grouper_string_sequences"]:
        types = seq["key_types"]
        if impl == "gpu":
            from arrow1_b200 import capi
            gb = capi.GroupBy(ctx, [TYPE_BY_NAME[t] for t in types], [], [])
            for b in seq["batches"]:
                keys = [_json_column(v, t).to_device() for v, t in zip(b["keys"], types)]
                ids = gb.grouper_consume(keys).cpu().numpy().view(np.uint32)
                np.testing.assert_array_equal(ids, np.array(b["ids"], dtype=np.uint32), err_msg=str((types, b)))
            gb.close()
            continue
        acc, expected = [[] for _ in types], []
        for b in seq["batches"]:
            for i, v in enumerate(b["keys"]):
                acc[i] += v
            expected += b["ids"]
            r = oracle.groupby([_json_column(v, t) for v, t in zip(acc, types)], [], [], want_ids=True, kind=impl)
            np.testing.assert_array_equal(r.ids, np.array(expected, dtype=np.uint32), err_msg=str((types, b)))


@pytest.mark.parametrize("impl", IMPLS[1:])
def test_group_by_string_key_golden(impl, request):
    """GroupBy.SumOnlyStringAndDictKeys (utf8 key): sums, the null sum of a group without values, group order, the
    null key as its own group."""
    e = load_golden("group_by.json")["sum_only_string_key"]
    ctx = _ctx(impl, request)
    r = run_groupby(impl, [Column.from_strings(e["key"])], [AggSpec(cols.HASH_SUM)], [column_from_json(e["argument"], "double")], ctx)
    assert r.num_groups == 4
    _cmp(r.aggs[0][0], e["hash_sum"], "double")
    vals, valid = r.keys[0]
    assert [v.decode() if ok else None for v, ok in zip(vals, valid)] == e["key_0"]


@pytest.mark.parametrize("impl", IMPLS)
def test_sort_indices_multi_key_golden(impl, request):
    """TestRecordBatchSortIndices / TestTableSortIndices: two sort keys, nulls and NaNs, one chunk and several."""
    g = load_golden("sort_indices_multi.json")
    ctx = _ctx(impl, request)
    for case in g["cases"]:
        keys = [column_from_json(v, t) for v, t in zip(case["columns"], case["types"])]
        want = np.array(case["expected"], dtype=np.uint64)
        for cuts in case["chunks"]:
            if impl == "gpu":
                from arrow1_b200 import capi
                if len(cuts) == 2:
                    got = capi.sort_indices_multi(ctx, [k.to_device() for k in keys], case["orders"])
                else:
                    pieces = [[column_from_json(v[a:b], t).to_device() for a, b in zip(cuts[:-1], cuts[1:])]
                              for v, t in zip(case["columns"], case["types"])]
                    got = capi.sort_indices_multi(ctx, pieces, case["orders"])
                got = got.cpu().numpy().view(np.uint64)
            elif impl == "reference":
                got = oracle.sort_indices_multi(keys, case["orders"], kind="reference", as_table=len(cuts) > 2, n_chunks=len(cuts) - 1)
            else:
                got = oracle.sort_indices_multi(keys, case["orders"])
            np.testing.assert_array_equal(got, want, err_msg=case["name"])


# ---- unique / value_counts / dictionary_encode: the reference's own known-answer tests (vector_hash_test.cc) ----------
_HASH_NUMERIC = ["uint8", "int8", "uint16", "int16", "uint32", "int32", "uint64", "int64", "float", "double"]


def _hash_column(case, type_name):
    col = _json_column(case["values"], type_name) if not case.get("offset") else None
    if col is None:
        off = case["offset"]
        if type_name == "binary":
            col = cols.Column.from_strings(case["values"], offset=off)
        else:
            col = column_from_json(case["values"], type_name, offset=off)
        col.length = case["length"]
        col.null_count = -1
    return col


def _as_list(vals, valid, type_name):
    if type_name == "binary":
        return [v.decode() if ok else None for v, ok in zip(vals, valid)]
    return [(float(v) if type_name in ("float", "double") else int(v)) if ok else None for v, ok in zip(vals, valid)]


def _host_pair(col):
    h = col.to_host()
    if h.type == cols.BINARY:
        s = h.strings()
        return [b"" if x is None else x for x in s], h.valid()
    return h.values(), h.valid()


@pytest.mark.parametrize("impl", IMPLS[1:])
def test_vector_hash_golden(impl, request):
    g = load_golden("vector_hash.json")
    ctx = _ctx(impl, request)
    if impl == "gpu":
        from arrow1_b200 import compute

    def types_of(case):
        return _HASH_NUMERIC if case["type"] == "NUMERIC" else [case["type"]]

    for case in g["unique"]:
        for t in types_of(case):
            col = _hash_column(case, t)
            if impl == "gpu":
                vals, valid = _host_pair(compute.Unique(col.to_device(), ctx))
            else:
                (vals, valid), _ = oracle.value_counts(col, False)
            assert _as_list(vals, valid, t) == case["expected"], ("unique", t, case)
    for case in g["value_counts"]:
        for t in types_of(case):
            col = _hash_column(case, t)
            if impl == "gpu":
                keys, counts = compute.ValueCounts(col.to_device(), ctx)
                vals, valid = _host_pair(keys)
                counts = counts.cpu().numpy()
            else:
                (vals, valid), counts = oracle.value_counts(col, True)
            assert _as_list(vals, valid, t) == case["expected"], ("value_counts", t, case)
            assert [int(c) for c in counts] == case["counts"], ("value_counts", t, case)
    for case in g["dictionary_encode"]:
        for t in types_of(case):
            col = _hash_column(case, t)
            if impl == "gpu":
                idx, dictionary = compute.DictionaryEncode(col.to_device(), case["encode_nulls"], ctx)
                hi = idx.to_host()
                ivals, ivalid = hi.values(), hi.valid()
                dvals, dvalid = _host_pair(dictionary)
            else:
                ivals, ivalid, (dvals, dvalid) = oracle.dictionary_encode(col, case["encode_nulls"])
            assert [int(v) if ok else None for v, ok in zip(ivals, ivalid)] == case["indices"], ("dictionary_encode", t, case)
            assert _as_list(dvals, dvalid, t) == case["dictionary"], ("dictionary_encode", t, case)
