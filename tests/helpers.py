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


def run_groupby(impl, keys, aggs, args, ctx=None, want_ids=False, expected_groups=0) -> GroupRes:
    if impl in ("port", "reference"):
        r = oracle.groupby(keys, aggs, args, want_ids=want_ids, kind=impl)
        return GroupRes(r.num_groups, r.aggs, r.keys, r.ids)
    from arrow1_b200 import capi
    dkeys = [k.to_device() for k in keys]
    dargs = [a.to_device() for a in args]
    out = capi.group_by(ctx, dkeys, aggs, dargs, expected_groups=expected_groups)
    res_aggs = [[(c.to_host().values().copy(), c.to_host().valid()) for c in a] for a in out.aggs]
    res_keys = [(c.to_host().values().copy(), c.to_host().valid()) for c in out.keys]
    ids = None
    if want_ids:
        gb = capi.GroupBy(ctx, [k.type for k in keys], [], [], expected_groups)
        ids = gb.grouper_consume(dkeys).cpu().numpy().view(np.uint32)
        gb.close()
    return GroupRes(out.num_groups, res_aggs, res_keys, ids)


def _is_float(arr):
    return arr.dtype.kind == "f"


def assert_group_equal(a: GroupRes, b: GroupRes, aggs, rtol=1e-12):
    """Bit-exact for keys, counts, min/max and integer sums; float sums/means within rtol
    (north_star: <= 1e-12 relative for float64)."""
    assert a.num_groups == b.num_groups
    for (ka, va), (kb, vb) in zip(a.keys, b.keys):
        np.testing.assert_array_equal(va, vb)
        np.testing.assert_array_equal(ka[va].view(np.uint8), kb[vb].view(np.uint8))
    for spec, ca, cb in zip(aggs, a.aggs, b.aggs):
        for (xa, va), (xb, vb) in zip(ca, cb):
            np.testing.assert_array_equal(va, vb, err_msg=spec.name())
            if spec.kind in (cols.HASH_SUM, cols.HASH_MEAN) and _is_float(xa):
                fa, fb = xa[va], xb[vb]
                nan_a, nan_b = np.isnan(fa), np.isnan(fb)
                np.testing.assert_array_equal(nan_a, nan_b)
                np.testing.assert_allclose(fa[~nan_a], fb[~nan_b], rtol=rtol, atol=0, err_msg=spec.name())
            else:
                np.testing.assert_array_equal(xa[va].view(np.uint8), xb[vb].view(np.uint8), err_msg=spec.name())
    if a.ids is not None and b.ids is not None:
        np.testing.assert_array_equal(a.ids, b.ids)
