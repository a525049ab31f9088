"""b200_filter_fused_noalloc — Filter for callers that must not allocate (no counterpart in the reference):
the plan lives in a static device array inside the library, the call makes no
cudaMalloc / cudaFree.  Checked through the C ABI against the existing entry point b200_filter_fused (caller scratch)
on the same inputs: bit-exact data, validity and length, on two calls in a row (the second one shorter, so that what the
first call left in the shared array would show if it were read), plus the size limit.

Collected last on purpose (file name).  Passes alone in a fresh process and after the
other tests."""
import ctypes as C

import numpy as np
import pytest
import torch

from arrow1_b200 import capi, columns as cols
from arrow1_b200.columns import Column
from gen import rand_column

pytestmark = pytest.mark.gpu


def _fused(ctx, entry, values, mask, null_sel, cap, scratch_ptr=None):
    """One call of `entry` into freshly zeroed caller-owned outputs -> (length, data bytes, validity bytes)."""
    bits = cols.BITS[values.type]
    data = torch.zeros((cap + 7) // 8 + 64 if bits == 1 else max(cap, 1) * (bits // 8) + 64, dtype=torch.uint8, device="cuda")
    validity = torch.zeros((cap + 7) // 8 + 64, dtype=torch.uint8, device="cuda")
    has_validity, _ = capi.filter_out_spec(values, mask, null_sel)
    v, m = values.desc(), mask.desc()
    out_len = C.c_int64(-1)
    args = [ctx._h, C.byref(v), C.byref(m), C.c_int(null_sel)]
    if scratch_ptr is not None:
        args.append(C.c_void_p(scratch_ptr))
    args += [C.c_void_p(data.data_ptr()), C.c_void_p(validity.data_ptr()) if has_validity else None, C.c_int64(cap), C.byref(out_len)]
    ctx.check(entry(*args))
    ctx.synchronize()
    n = int(out_len.value)
    nb = (n + 7) // 8 if bits == 1 else n * (bits // 8)
    return n, data[:nb].cpu().numpy().tobytes(), validity[: (n + 7) // 8].cpu().numpy().tobytes() if has_validity else None


@pytest.mark.parametrize("t,null_sel", [(cols.INT64, 0), (cols.INT32, 1), (cols.BOOL, 0), (cols.UINT8, 1), (cols.DOUBLE, 0)])
def test_filter_fused_noalloc_matches_filter_fused_twice_in_a_row(ctx, t, null_sel):
    lib = capi.lib()
    lib.b200_filter_fused_noalloc_max_rows.restype = C.c_int64
    assert lib.b200_filter_fused_noalloc_max_rows() == 1 << 28
    rng = np.random.default_rng(1000 + t)
    # two calls in a row on the shared array: a long mask first, then a short one with other contents
    for n, p_true, nulls in ((1_300_007, 0.5, 0.1), (70_001, 0.9, 0.3), (0, 0.5, 0.0), (4097, 0.0, 0.0)):
        values = rand_column(rng, t, n, nulls).to_device()
        mask = Column.from_numpy(rng.random(n) < p_true, rng.random(n) > nulls if nulls else None).to_device()
        scratch = capi.FusedFilterScratch(ctx, max(n, 1))
        want = _fused(ctx, lib.b200_filter_fused, values, mask, null_sel, n, scratch.ptr)
        got = _fused(ctx, lib.b200_filter_fused_noalloc, values, mask, null_sel, n)
        assert got[0] == want[0]
        assert got[1] == want[1], "data differs at n=%d" % n
        assert got[2] == want[2], "validity differs at n=%d" % n


def test_filter_fused_noalloc_rejects_masks_beyond_its_limit(ctx):
    lib = capi.lib()
    lib.b200_filter_fused_noalloc_max_rows.restype = C.c_int64
    too_long = lib.b200_filter_fused_noalloc_max_rows() + 1
    buf = torch.zeros(4096, dtype=torch.uint8, device="cuda")
    # descriptors only: the call must refuse before anything is launched or read
    values = Column(cols.UINT8, buf, None, too_long, 0, 0)
    mask = Column(cols.BOOL, buf, None, too_long, 0, 0)
    v, m = values.desc(), mask.desc()
    out_len = C.c_int64(-1)
    code = lib.b200_filter_fused_noalloc(ctx._h, C.byref(v), C.byref(m), C.c_int(0), C.c_void_p(buf.data_ptr()), None,
                                         C.c_int64(16), C.byref(out_len))
    assert code == 6, "CapacityError expected, got %d" % code   # arrow::StatusCode::CapacityError
    with pytest.raises(capi.B200Error) as e:
        ctx.check(code)
    assert "maximum" in str(e.value)
    ctx.synchronize()
    # and the entry point still works afterwards
    rng = np.random.default_rng(5)
    n = 10_000
    values = rand_column(rng, cols.INT64, n, 0.0).to_device()
    mask = Column.from_numpy(rng.random(n) < 0.5).to_device()
    scratch = capi.FusedFilterScratch(ctx, n)
    assert _fused(ctx, lib.b200_filter_fused_noalloc, values, mask, 0, n) == _fused(ctx, lib.b200_filter_fused, values, mask, 0, n, scratch.ptr)
