"""Filter parity: CUDA path (through the C ABI) vs the oracle, bit-exact including the bytes
under null slots, the validity bitmap padding, null_count and the allocate-validity rule."""
import numpy as np
import pytest

import oracle
from arrow1_b200 import columns as cols
from arrow1_b200.columns import Column
from conftest import rand_u64
from gen import ALL_NUMERIC, rand_column, rand_mask
from helpers import assert_filter_equal, run_filter

pytestmark = pytest.mark.gpu
ORACLE = "reference" if oracle.have_reference() else "port"


@pytest.mark.parametrize("t", [cols.BOOL] + ALL_NUMERIC)
def test_filter_random_matrix(ctx, t):
    # vector_selection_test.cc:1618-1636
    rng = np.random.default_rng(4321 + t)
    for null_prob in (0.0, 0.01, 0.1, 0.999, 1.0):
        for true_prob in (0.0, 0.1, 0.999, 1.0):
            for (voff, moff) in ((0, 0), (3, 4)):
                for mask_null in (0.0, 0.05):
                    values = rand_column(rng, t, 1024, null_prob, offset=voff)
                    mask = rand_mask(rng, 1024, true_prob, mask_null, offset=moff)
                    for null_sel in (0, 1):
                        a = run_filter("gpu", values, mask, null_sel, ctx)
                        assert_filter_equal(run_filter("gpu_fused", values, mask, null_sel, ctx), a)
                        b = run_filter(ORACLE, values, mask, null_sel)
                        assert_filter_equal(a, b)


@pytest.mark.parametrize("n", [1, 31, 32, 33, 63, 64, 65, 511, 512, 513, 4095, 4096, 4097, 262143, 262144, 262145,
                               1000003])
def test_filter_sizes_and_offsets(ctx, n):
    rng = np.random.default_rng(n)
    for t in (cols.INT64, cols.UINT8, cols.BOOL, cols.INT32, cols.INT16):
        for (voff, moff) in ((0, 0), (1, 7), (13, 64), (65, 3)):
            values = rand_column(rng, t, n, 0.1, offset=voff)
            mask = rand_mask(rng, n, 0.5, 0.05, offset=moff)
            for null_sel in (0, 1):
                assert_filter_equal(run_filter("gpu_fused", values, mask, null_sel, ctx), run_filter(ORACLE, values, mask, null_sel))
                assert_filter_equal(run_filter("gpu", values, mask, null_sel, ctx),
                                    run_filter(ORACLE, values, mask, null_sel))


def test_filter_unknown_null_count(ctx):
    # vector_selection_test.cc:445-455
    rng = np.random.default_rng(7)
    values = rand_column(rng, cols.INT32, 5000, 0.2)
    mask = rand_mask(rng, 5000, 0.5, 0.1)
    values.null_count = -1
    mask.null_count = -1
    for null_sel in (0, 1):
        assert_filter_equal(run_filter("gpu", values, mask, null_sel, ctx), run_filter(ORACLE, values, mask, null_sel))


def test_filter_count_matches(ctx):
    from arrow1_b200 import capi
    rng = np.random.default_rng(3)
    for n in (0, 1, 100, 4096, 300000):
        mask = rand_mask(rng, n, 0.3, 0.2, offset=5)
        for null_sel in (0, 1):
            exp = run_filter(ORACLE, rand_column(rng, cols.UINT8, n), mask, null_sel).length
            assert capi.filter_count(ctx, mask.to_device(), null_sel) == exp


def test_filter_non_boolean_mask_not_implemented(ctx):
    from arrow1_b200 import capi
    v = Column.from_numpy(np.arange(4, dtype=np.int32)).to_device()
    with pytest.raises(capi.B200Error) as ei:
        capi.filter(ctx, v, v, 0)
    assert ei.value.code == 10  # NotImplemented (vector_selection.cc:1885-1887)


def test_filter_plan_reused_for_several_columns(ctx):
    # FilterRecordBatch semantics == per-column array_filter (vector_selection_test.cc:208-236)
    from arrow1_b200 import capi
    rng = np.random.default_rng(17)
    n = 100000
    mask = rand_mask(rng, n, 0.96, 0.01)
    plan = capi.FilterPlan(ctx, mask.to_device(), 1, sync=True)
    for t in (cols.DOUBLE, cols.UINT8, cols.INT32):
        values = rand_column(rng, t, n, 0.02)
        out = plan.apply(values.to_device())
        ref = run_filter(ORACLE, values, mask, 1)
        assert out.length == ref.length
        assert out.data_bytes() == ref.data_bytes
        assert out.validity_bytes() == ref.validity_bytes


def test_filter_async_plan_keeps_length_on_device(ctx):
    from arrow1_b200 import capi
    rng = np.random.default_rng(19)
    n = 50000
    mask = rand_mask(rng, n, 0.5)
    values = rand_column(rng, cols.INT64, n, 0.01)
    plan = capi.FilterPlan(ctx, mask.to_device(), 0, sync=False)
    out = plan.apply(values.to_device())
    ctx.synchronize()
    length = int(plan.length_tensor().cpu()[0])
    ref = run_filter(ORACLE, values, mask, 0)
    assert length == ref.length
    assert out.col.data[: length * 8].cpu().numpy().tobytes() == ref.data_bytes
    assert out.col.validity[: (length + 7) // 8].cpu().numpy().tobytes() == ref.validity_bytes


def test_filter_config1_full_size(ctx):
    """BASELINE config 1: 10 M int64, 50 % mask, 1 % nulls, bit-exact vs the oracle."""
    n = 10_000_000
    with np.errstate(over="ignore"):
        values = rand_u64(1, n).view(np.int64)
        valid = (rand_u64(2, n) % np.uint64(100)) != 0
        bits = (rand_u64(3, n) & np.uint64(1)).astype(np.bool_)
    v = Column.from_numpy(values, valid)
    m = Column.from_numpy(bits)
    assert_filter_equal(run_filter("gpu_fused", v, m, 0, ctx), run_filter("gpu", v, m, 0, ctx))
    a = run_filter("gpu", v, m, 0, ctx)
    b = run_filter(ORACLE, v, m, 0)
    assert_filter_equal(a, b)
    # size-independent properties: length == popcount(mask); values keep their order
    assert a.length == int(bits.sum())
    np.testing.assert_array_equal(a.values, values[bits])


def test_filter_idempotent_with_all_true_mask(ctx):
    rng = np.random.default_rng(23)
    n = 123457
    values = rand_column(rng, cols.DOUBLE, n, 0.3, offset=9)
    mask = Column.from_numpy(np.ones(n, dtype=np.bool_))
    a = run_filter("gpu", values, mask, 0, ctx)
    assert a.length == n
    np.testing.assert_array_equal(a.valid, values.valid())
    np.testing.assert_array_equal(a.values.view(np.uint64), values.values().view(np.uint64))


def test_filter_apply_rejects_small_output_and_sync_reports_it(ctx):
    """an output smaller than the plan's length is never written past; the context reports it at the next
    synchronisation (B200_UNKNOWN_ERROR), once.  Same channel as the sort's look-back bail-out (injected here)."""
    import ctypes as C
    import torch
    from arrow1_b200 import capi
    from arrow1_b200.capi import B200Error
    rng = np.random.default_rng(3)
    n = 100_000
    values = Column.from_numpy(rng.integers(0, 1 << 40, size=n, dtype=np.int64)).to_device()
    mask = Column.from_numpy(rng.random(n) < 0.5).to_device()
    plan = capi.FilterPlan(ctx, mask, 0, sync=True)
    assert plan.out_length > 1000
    small = 1000
    out = torch.full((small * 8 + 4096,), 0x5A, dtype=torch.uint8, device="cuda")
    v, m = values.desc(), mask.desc()
    ctx.check(capi.lib().b200_filter_apply(ctx._h, C.byref(v), C.byref(m), C.c_int(0), C.c_void_p(plan.buf.data_ptr()),
                                           C.c_void_p(out.data_ptr()), None, C.c_int64(small)))
    with pytest.raises(B200Error) as e:
        ctx.synchronize()
    assert e.value.code == 9 and "filter output" in str(e.value)
    assert bool((out == 0x5A).all()), "nothing may be written when the output is too small"
    ctx.synchronize()   # the flag is cleared by the report
    ctx.check(capi.lib().b200_ctx_inject_async_error(ctx._h, C.c_uint32(1)))
    with pytest.raises(B200Error) as e:
        ctx.synchronize()
    assert "look-back" in str(e.value)
    ctx.synchronize()


def test_filter_fused_async_and_small_capacity(ctx):
    """b200_filter_fused without a host round trip: the length stays on the device; a capacity below the output
    length writes nothing past it and is reported at the next synchronisation."""
    import torch
    from arrow1_b200 import capi
    from arrow1_b200.capi import B200Error
    rng = np.random.default_rng(31)
    n = 1_000_003
    values = Column.from_numpy(rng.integers(-2**62, 2**62, size=n, dtype=np.int64), rng.random(n) > 0.01)
    mask = Column.from_numpy(rng.random(n) < 0.5)
    want = run_filter(ORACLE, values, mask, 0)
    dv, dm = values.to_device(), mask.to_device()
    scratch = capi.FusedFilterScratch(ctx, n)
    for _ in range(3):   # the scratch is reused call after call
        out = capi.filter_fused(ctx, dv, dm, 0, scratch=scratch, sync=False)
        ctx.synchronize()
        m = int(scratch.length_tensor().item())
        assert m == want.length
        assert out.col.data[: m * 8].cpu().numpy().tobytes() == want.data_bytes
        assert out.col.validity[: (m + 7) // 8].cpu().numpy().tobytes() == want.validity_bytes
    small = want.length - 5000
    data = torch.full((small * 8 + 65536,), 0x5A, dtype=torch.uint8, device="cuda")
    capi.filter_fused(ctx, dv, dm, 0, out_data=data, capacity=small, scratch=scratch, sync=False)
    with pytest.raises(B200Error) as e:
        ctx.synchronize()
    assert "filter output" in str(e.value)
    assert bool((data[small * 8:] == 0x5A).all())
    ctx.synchronize()
