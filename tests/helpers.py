"""Implementation-agnostic runners: the same case can be pushed through the oracle port, the
unmodified reference (when built) or the CUDA path through the C ABI, and compared."""
from __future__ import annotations

import numpy as np

import oracle
from arrow1_b200 import columns as cols
from arrow1_b200.columns import Column


class FilterRes:
    def __init__(self, length, null_count, has_validity, values, valid, data_bytes, validity_bytes):
        self.length, self.null_count, self.has_validity = length, null_count, has_validity
        self.values, self.valid = values, valid
        self.data_bytes, self.validity_bytes = data_bytes, validity_bytes


def run_filter(impl, values: Column, mask: Column, null_selection: int, ctx=None) -> FilterRes:
    if impl in ("port", "reference"):
        r = oracle.filter(values, mask, null_selection, kind=impl)
        return FilterRes(r.length, r.null_count, r.has_validity, r.values(), r.valid(), r.data_bytes(),
                         r.validity_bytes())
    from arrow1_b200 import capi
    if impl == "gpu_fused":   # one launch, output sized by the mask length, no host round trip inside
        out = capi.filter_fused(ctx, values.to_device(), mask.to_device(), null_selection)
    else:
        out = capi.filter(ctx, values.to_device(), mask.to_device(), null_selection)
    col = out.col
    host = col.to_host()
    valid = host.valid() if out.has_validity else np.ones(col.length, dtype=np.bool_)
    return FilterRes(col.length, col.null_count, out.has_validity, host.values().copy(), valid, out.data_bytes(),
                     out.validity_bytes())


def assert_filter_equal(a: FilterRes, b: FilterRes, bit_exact=True):
    assert a.length == b.length
    assert a.null_count == b.null_count
    assert a.has_validity == b.has_validity
    if bit_exact:
        # data bytes under nulls included (vector_selection.cc:624-629, 743-746); bitmap padding zero
        assert a.data_bytes == b.data_bytes
        assert a.validity_bytes == b.validity_bytes
    else:
        np.testing.assert_array_equal(a.valid, b.valid)
        np.testing.assert_array_equal(a.values[a.valid].view(np.uint8), b.values[b.valid].view(np.uint8))


def run_sort(impl, values: Column, order: int, index_base: int = 0, ctx=None) -> np.ndarray:
    if impl in ("port", "reference"):
        return oracle.sort_indices(values, order, index_base, kind=impl)
    from arrow1_b200 import capi
    out = capi.sort_indices(ctx, values.to_device(), order, index_base)
    return out.cpu().numpy().view(np.uint64)


class GroupRes:
    def __init__(self, num_groups, aggs, keys, ids=None):
        self.num_groups, self.aggs, self.keys, self.ids = num_groups, aggs, keys, ids


def run_groupby(impl, keys, aggs, args, ctx=None, want_ids=False, expected_groups=0, options=None) -> GroupRes:
    if impl in ("port", "reference"):
        r = oracle.groupby(keys, aggs, args, want_ids=want_ids, kind=impl)
        return GroupRes(r.num_groups, r.aggs, r.keys, r.ids)
    from arrow1_b200 import capi
    dev = {}   # one device copy per distinct host column (aggregates sharing a column share its buffers)
    for c in list(keys) + list(args):
        if id(c) not in dev:
            dev[id(c)] = c.to_device()
    dkeys = [dev[id(k)] for k in keys]
    dargs = [dev[id(a)] for a in args]
    out = capi.group_by(ctx, dkeys, aggs, dargs, expected_groups=expected_groups, options=options)
    res_aggs = [[(c.to_host().values().copy(), c.to_host().valid()) for c in a] for a in out.aggs]
    res_keys = [_host_key(c) for c in out.keys]
    ids = None
    if want_ids:
        gb = capi.GroupBy(ctx, [k.type for k in keys], [], [], expected_groups)
        ids = gb.grouper_consume(dkeys).cpu().numpy().view(np.uint32)
        gb.close()
    return GroupRes(out.num_groups, res_aggs, res_keys, ids)


def _host_key(c):
    """(values, valid) of a key column; a BINARY column becomes an object array of bytes (b"" where null)."""
    h = c.to_host()
    if h.type == cols.BINARY:
        vals = np.empty(h.length, dtype=object)
        for i, x in enumerate(h.strings()):
            vals[i] = b"" if x is None else x
        return vals, h.valid()
    return h.values().copy(), h.valid()


def _is_float(arr):
    return arr.dtype.kind == "f"


def abs_sum_scales(keys, aggs, args, oracle_kind="port"):
    """Per-group sum of |v| for every float sum/mean aggregate: the scale the float tolerance is
    relative to (a reordered sum cannot be relatively accurate against a cancelled result)."""
    scales = {}
    for j, (spec, arg) in enumerate(zip(aggs, args)):
        if spec.kind in (cols.HASH_SUM, cols.HASH_MEAN) and arg.type in (cols.FLOAT, cols.DOUBLE):
            v = np.abs(np.nan_to_num(arg.values().astype(np.float64), nan=0.0, posinf=0.0, neginf=0.0))
            absarg = Column.from_numpy(v, arg.valid() if arg.validity is not None else None)
            r = oracle.groupby(keys, [cols.AggSpec(spec.kind)], [absarg], kind=oracle_kind)
            scales[j] = r.aggs[0][0][0]
    return scales


def assert_group_equal(a: GroupRes, b: GroupRes, aggs, rtol=1e-12, scales=None):
    """Bit-exact for keys, counts, min/max and integer sums.  Float sums / means: |gpu - ref| <=
    rtol * S with S = the group's sum (mean) of |v| — i.e. a relative tolerance of 1e-12 for
    same-sign data such as the benchmark columns (north_star: <= 1e-12 relative for float64);
    without `scales` S = |ref|."""
    assert a.num_groups == b.num_groups
    for (ka, va), (kb, vb) in zip(a.keys, b.keys):
        np.testing.assert_array_equal(va, vb)
        if ka.dtype == object:   # variable-length keys: bytes objects
            assert list(ka[va]) == list(kb[vb])
        else:
            np.testing.assert_array_equal(ka[va].view(np.uint8), kb[vb].view(np.uint8))
    for spec, ca, cb in zip(aggs, a.aggs, b.aggs):
        for (xa, va), (xb, vb) in zip(ca, cb):
            np.testing.assert_array_equal(va, vb, err_msg=spec.name())
            if spec.kind in (cols.HASH_SUM, cols.HASH_MEAN) and _is_float(xa):
                fa, fb = xa[va], xb[vb]
                nan_a, nan_b = np.isnan(fa), np.isnan(fb)
                np.testing.assert_array_equal(nan_a, nan_b)
                j = list(aggs).index(spec) if scales is None else None
                scale = np.abs(fb)
                if scales is not None:
                    idx = [i for i, s_ in enumerate(aggs) if s_ is spec][0]
                    if idx in scales:
                        scale = np.maximum(scale, scales[idx][vb])
                fin = ~nan_a & np.isfinite(fb)
                np.testing.assert_array_equal(fa[~nan_a & ~fin], fb[~nan_b & ~fin])
                err = np.abs(fa[fin] - fb[fin])
                assert (err <= rtol * scale[fin]).all(), "%s: max err/scale %g" % (
                    spec.name(), float(np.max(err / np.maximum(scale[fin], 1e-300))))
            else:
                np.testing.assert_array_equal(xa[va].view(np.uint8), xb[vb].view(np.uint8), err_msg=spec.name())
    if a.ids is not None and b.ids is not None:
        np.testing.assert_array_equal(a.ids, b.ids)


def MULTI_KEY_GOLDEN():
    """tests/golden/sort_indices_multi.json (vector_sort_test.cc:763-842, 886-980) as (keys, orders, expected)."""
    from conftest import column_from_json, load_golden
    for case in load_golden("sort_indices_multi.json")["cases"]:
        keys = [column_from_json(v, t) for v, t in zip(case["columns"], case["types"])]
        yield keys, case["orders"], case["expected"]


def multi_key_cases(rng, n):
    import numpy as np
    from arrow1_b200 import columns as cols
    from gen import rand_column
    """(keys, orders) lists: few distinct values per key so that ties reach the later keys; nulls, NaNs, -0.0."""
    def small(t, k, nulls):
        c = rand_column(rng, t, n, nulls, lo=0, hi=k) if t not in (cols.FLOAT, cols.DOUBLE) else None
        if c is None:
            v = rng.integers(-k, k, size=n).astype(np.float32 if t == cols.FLOAT else np.float64) / 2
            v[rng.random(n) < 0.1] = np.nan
            v[rng.random(n) < 0.05] = -0.0
            c = cols.Column.from_numpy(v, (rng.random(n) >= nulls) if nulls else None)
        return c
    yield [small(cols.INT32, 5, 0.1), small(cols.DOUBLE, 4, 0.1)], [0, 1]
    yield [small(cols.UINT8, 3, 0.0), small(cols.INT64, 3, 0.2), small(cols.FLOAT, 3, 0.0)], [1, 0, 0]
    yield [small(cols.DOUBLE, 2, 0.3), small(cols.DOUBLE, 2, 0.3), small(cols.INT16, 2, 0.3), small(cols.UINT64, 50, 0.0)], [1, 1, 0, 1]
    yield [small(cols.INT8, 4, 0.0), small(cols.UINT16, 4, 0.0)], [0, 0]
