"""ctypes binding of the C ABI (include/arrow1_b200.h) over torch-owned device memory.

PyTorch is used here for device buffers and streams only; every computation goes through
libarrow1_b200.so.  There is no fallback: a missing library or GPU raises.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence

import numpy as np
import torch

from .columns import (BINARY, BITS, BOOL, DOUBLE, HASH_MIN_MAX, INT64, NP_OF, UINT64, AggSpec, Column, agg_out_types,
                      b200_agg_spec, b200_col, b200_out_col, col_array)

_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libarrow1_b200.so")
_lib = None

STATUS_NAMES = {0: "OK", 1: "OutOfMemory", 2: "KeyError", 3: "TypeError", 4: "Invalid", 5: "IOError",
                6: "CapacityError", 7: "IndexError", 9: "UnknownError", 10: "NotImplemented"}


class B200Error(RuntimeError):
    """Carries the arrow::StatusCode-valued status of a failed C-ABI call."""

    def __init__(self, code: int, msg: str):
        super().__init__("%s: %s" % (STATUS_NAMES.get(code, str(code)), msg))
        self.code = code
        self.message = msg


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        path = os.environ.get("B200_LIB", _LIB_PATH)   # tuning sweeps load alternative builds
        if not os.path.exists(path):
            raise RuntimeError("libarrow1_b200.so is not built: run `python -m arrow1_b200.build` "
                               "(there is no CPU fallback)")
        L = C.CDLL(path)
        L.b200_last_error.restype = C.c_char_p
        L.b200_version.restype = C.c_char_p
        L.b200_ctx_launch_count.restype = C.c_int64
        L.b200_filter_plan_length_ptr.restype = C.c_void_p
        L.b200_filter_plan_length_ptr.argtypes = [C.c_void_p]
        _lib = L
    return _lib


class Context:
    """One device + one stream (b200_ctx)."""

    def __init__(self, device: int = 0, stream: Optional[torch.cuda.Stream] = None):
        if not torch.cuda.is_available():
            raise RuntimeError("arrow1_b200 needs a CUDA device (sm_100a); no CPU fallback exists")
        self.device = device
        torch.cuda.set_device(device)
        self.stream = stream if stream is not None else torch.cuda.current_stream(device)
        self._h = C.c_void_p()
        code = lib().b200_ctx_create(C.c_int(device), C.c_void_p(self.stream.cuda_stream), C.byref(self._h))
        if code != 0:
            raise B200Error(code, "b200_ctx_create failed")

    def close(self):
        if self._h:
            lib().b200_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def check(self, code: int):
        if code != 0:
            raise B200Error(code, lib().b200_last_error(self._h).decode())

    def synchronize(self):
        self.check(lib().b200_ctx_synchronize(self._h))

    def set_option(self, name: str, value: int):
        self.check(lib().b200_ctx_set_option(self._h, name.encode(), C.c_int64(value)))

    @property
    def launches(self) -> int:
        return int(lib().b200_ctx_launch_count(self._h))

    def empty(self, nbytes: int) -> torch.Tensor:
        # rounded up so bitmap tails and vector stores stay inside the allocation
        return torch.empty(max((nbytes + 63) // 64 * 64, 64), dtype=torch.uint8, device="cuda:%d" % self.device)


# ---------------------------------------------------------------------------------------------
# Filter


class FilterOutput:
    def __init__(self, col: Column, has_validity: bool):
        self.col = col
        self.has_validity = has_validity

    @property
    def length(self):
        return self.col.length

    @property
    def null_count(self):
        return self.col.null_count

    def data_bytes(self) -> bytes:
        bits = BITS[self.col.type]
        n = (self.col.length + 7) // 8 if bits == 1 else self.col.length * (bits // 8)
        return self.col.data[:n].cpu().numpy().tobytes()

    def validity_bytes(self) -> Optional[bytes]:
        if not self.has_validity:
            return None
        return self.col.validity[: (self.col.length + 7) // 8].cpu().numpy().tobytes()


def filter_out_spec(values: Column, mask: Column, null_selection: int):
    av, nc = C.c_int(0), C.c_int64(0)
    v, m = values.desc(), mask.desc()
    lib().b200_filter_out_spec(C.byref(v), C.byref(m), C.c_int(null_selection), C.byref(av), C.byref(nc))
    return bool(av.value), int(nc.value)


class FilterPlan:
    """Device-side scan of one mask, reusable for several value columns (b200_filter_plan)."""

    def __init__(self, ctx: Context, mask: Column, null_selection: int = 0, sync: bool = True):
        self.ctx, self.mask, self.null_selection = ctx, mask, null_selection
        nbytes = C.c_size_t(0)
        ctx.check(lib().b200_filter_plan_bytes(C.c_int64(mask.length), C.byref(nbytes)))
        self.buf = ctx.empty(nbytes.value)
        m = mask.desc()
        out_len = C.c_int64(-1)
        ctx.check(lib().b200_filter_plan(ctx._h, C.byref(m), C.c_int(null_selection), C.c_void_p(self.buf.data_ptr()),
                                         C.byref(out_len) if sync else None))
        self.out_length = int(out_len.value) if sync else None

    def length_tensor(self) -> torch.Tensor:
        return self.buf[:8].view(torch.int64)

    def apply(self, values: Column, out_data: Optional[torch.Tensor] = None,
              out_validity: Optional[torch.Tensor] = None) -> FilterOutput:
        ctx = self.ctx
        cap = self.out_length if self.out_length is not None else self.mask.length
        has_validity, null_count = filter_out_spec(values, self.mask, self.null_selection)
        bits = BITS[values.type]
        if out_data is None:
            out_data = ctx.empty((cap + 7) // 8 if bits == 1 else cap * (bits // 8))
        if has_validity and out_validity is None:
            out_validity = ctx.empty((cap + 7) // 8)
        v, m = values.desc(), self.mask.desc()
        ctx.check(lib().b200_filter_apply(ctx._h, C.byref(v), C.byref(m), C.c_int(self.null_selection),
                                          C.c_void_p(self.buf.data_ptr()), C.c_void_p(out_data.data_ptr()),
                                          C.c_void_p(out_validity.data_ptr()) if has_validity else None,
                                          C.c_int64(cap)))
        col = Column(values.type, out_data, out_validity if has_validity else None,
                     self.out_length if self.out_length is not None else -1, 0, null_count)
        return FilterOutput(col, has_validity)


class FusedFilterScratch:
    """Look-back state of b200_filter_fused for masks of up to `length` rows; its first 8 bytes hold the int64
    output length of the last call (device)."""

    def __init__(self, ctx: Context, length: int):
        nbytes = C.c_size_t(0)
        ctx.check(lib().b200_filter_fused_scratch_bytes(C.c_int64(length), C.byref(nbytes)))
        self.length = length
        self.buf = ctx.empty(nbytes.value + 256)
        self.ptr = (self.buf.data_ptr() + 255) // 256 * 256

    def length_tensor(self) -> torch.Tensor:
        off = self.ptr - self.buf.data_ptr()
        return self.buf[off: off + 8].view(torch.int64)


def filter_fused(ctx: Context, values: Column, mask: Column, null_selection: int = 0,
                 out_data: Optional[torch.Tensor] = None, out_validity: Optional[torch.Tensor] = None,
                 capacity: Optional[int] = None, scratch: Optional[FusedFilterScratch] = None, sync: bool = True) -> FilterOutput:
    """b200_filter_fused: Filter in one launch into an output of `capacity` rows (default: the mask length).
    sync=False leaves the output length on the device (scratch.length_tensor()); the Column's length is then -1."""
    cap = mask.length if capacity is None else capacity
    has_validity, null_count = filter_out_spec(values, mask, null_selection)
    bits = BITS[values.type]
    if out_data is None:
        out_data = ctx.empty((cap + 7) // 8 + 8 if bits == 1 else max(cap, 1) * (bits // 8))
    if has_validity and out_validity is None:
        out_validity = ctx.empty((cap + 7) // 8 + 8)
    if scratch is None or scratch.length < mask.length:
        scratch = FusedFilterScratch(ctx, mask.length)
    v, m = values.desc(), mask.desc()
    out_len = C.c_int64(-1)
    ctx.check(lib().b200_filter_fused(ctx._h, C.byref(v), C.byref(m), C.c_int(null_selection), C.c_void_p(scratch.ptr),
                                      C.c_void_p(out_data.data_ptr()), C.c_void_p(out_validity.data_ptr()) if has_validity else None,
                                      C.c_int64(cap), C.byref(out_len) if sync else None))
    col = Column(values.type, out_data, out_validity if has_validity else None, int(out_len.value) if sync else -1, 0, null_count)
    out = FilterOutput(col, has_validity)
    out.scratch = scratch
    return out


CMP_OPS = {"equal": 0, "not_equal": 1, "greater": 2, "greater_equal": 3, "less": 4, "less_equal": 5}


def compare_scalar(ctx: Context, values: Column, op: str, scalar) -> Column:
    """== CallFunction(op, {values, scalar}) of compute/kernels/scalar_compare.cc on a device column:
    boolean Column whose validity is the input's (shared buffer)."""
    n = values.length
    out = ctx.empty((n + 31) // 32 * 4 + 8)
    sc = np.array([scalar], dtype=NP_OF[values.type])
    v = values.desc()
    ctx.check(lib().b200_compare_scalar(ctx._h, C.byref(v), C.c_int(CMP_OPS[op]), C.c_void_p(sc.ctypes.data),
                                        C.c_void_p(out.data_ptr())))
    if values.validity is not None and values.offset != 0:
        raise B200Error(10, "compare_scalar: sliced nullable input (validity cannot be shared at offset 0)")
    return Column(BOOL, out, values.validity, n, 0, values.null_count if values.validity is not None else 0)


def take(ctx: Context, values, indices, boundscheck: bool = True, want_null_count: bool = True) -> Column:
    """== compute::Take(values, indices, TakeOptions(boundscheck)) on device columns.  `values` is a Column or a
    sequence of Columns of one type (the chunks of a ChunkedArray: indices address their concatenation,
    b200_take_chunked).  `indices` is a Column of any integer type, or the int64-typed tensor b200_sort_indices
    returns (uint64 indices, no nulls)."""
    from .columns import UINT64
    if isinstance(indices, torch.Tensor):
        indices = Column(UINT64, indices.view(torch.uint8).reshape(-1), None, indices.numel(), 0, 0)
    chunks = None if isinstance(values, Column) else list(values)
    vtype = values.type if chunks is None else chunks[0].type
    n = indices.length
    bits = BITS[vtype]
    out = ctx.empty((n + 31) // 32 * 4 + 8 if bits == 1 else max(n, 1) * (bits // 8))
    out_v = ctx.empty((n + 31) // 32 * 4 + 8)
    nc = C.c_int64(-1)
    i = indices.desc()
    tail = (C.byref(i), C.c_int(1 if boundscheck else 0), C.c_void_p(out.data_ptr()), C.c_void_p(out_v.data_ptr()),
            C.byref(nc) if want_null_count else None)
    if chunks is None:
        v = values.desc()
        ctx.check(lib().b200_take(ctx._h, C.byref(v), *tail))
    else:
        ctx.check(lib().b200_take_chunked(ctx._h, col_array(chunks), C.c_int(len(chunks)), *tail))
    return Column(vtype, out, out_v, n, 0, int(nc.value) if want_null_count else -1)


ARITH_OPS = {"add": 0, "subtract": 1, "multiply": 2}


def arithmetic(ctx: Context, op: str, a, b) -> Column:
    """== CallFunction(op, {a, b}) for "add" / "subtract" / "multiply" (compute/kernels/scalar_arithmetic.cc) where
    a and b are device Columns of one numeric type or Python / numpy scalars (at least one Column).  The result's
    validity is the AND of the operands' (null_count -1 = unknown when any operand may hold nulls)."""
    ca = a if isinstance(a, Column) else None
    cb = b if isinstance(b, Column) else None
    c = ca if ca is not None else cb
    if c is None:
        raise B200Error(4, "arithmetic: at least one operand must be a column")
    n = c.length
    width = BITS[c.type] // 8
    out = ctx.empty(max(n, 1) * max(width, 1))
    nullable = any(x is not None and x.validity is not None and x.null_count != 0 for x in (ca, cb))
    out_v = ctx.empty((n + 31) // 32 * 4 + 8) if nullable else None
    sa = None if ca is not None else np.array([a], dtype=NP_OF[c.type])
    sb = None if cb is not None else np.array([b], dtype=NP_OF[c.type])
    da = ca.desc() if ca is not None else None
    db = cb.desc() if cb is not None else None
    ctx.check(lib().b200_arithmetic(ctx._h, C.c_int(ARITH_OPS[op]),
                                    C.byref(da) if da is not None else None, C.c_void_p(sa.ctypes.data) if sa is not None else None,
                                    C.byref(db) if db is not None else None, C.c_void_p(sb.ctypes.data) if sb is not None else None,
                                    C.c_void_p(out.data_ptr()), C.c_void_p(out_v.data_ptr()) if nullable else None))
    return Column(c.type, out, out_v, n, 0, -1 if nullable else 0)


def filter_count(ctx: Context, mask: Column, null_selection: int = 0) -> int:
    m = mask.desc()
    n = C.c_int64(0)
    ctx.check(lib().b200_filter_count(ctx._h, C.byref(m), C.c_int(null_selection), C.byref(n)))
    return int(n.value)


def filter(ctx: Context, values: Column, mask: Column, null_selection: int = 0) -> FilterOutput:
    """== compute::Filter(values, mask, FilterOptions(null_selection)) on device columns."""
    if values.length != mask.length:
        # ExecBatchIterator::Make (compute/exec.cc:186-189)
        raise B200Error(4, "Array arguments must all be the same length")
    return FilterPlan(ctx, mask, null_selection, sync=True).apply(values)


# ---------------------------------------------------------------------------------------------
# SortIndices


def sort_scratch_bytes(ctx: Context, length: int) -> int:
    n = C.c_size_t(0)
    ctx.check(lib().b200_sort_indices_scratch_bytes(C.c_int64(length), C.byref(n)))
    return int(n.value)


def sort_indices(ctx: Context, values: Column, order: int = 0, index_base: int = 0,
                 out: Optional[torch.Tensor] = None, scratch: Optional[torch.Tensor] = None) -> torch.Tensor:
    """== compute::SortIndices(values, order): int64-typed tensor holding the uint64 indices."""
    n = values.length
    if out is None:
        out = torch.empty(max(n, 1), dtype=torch.int64, device="cuda:%d" % ctx.device)
    if scratch is None:
        scratch = ctx.empty(sort_scratch_bytes(ctx, n))
    v = values.desc()
    ctx.check(lib().b200_sort_indices(ctx._h, C.byref(v), C.c_int(order), C.c_uint64(index_base),
                                      C.c_void_p(out.data_ptr()), C.c_void_p(scratch.data_ptr()),
                                      C.c_size_t(scratch.numel())))
    return out[:n]


def sort_indices_plan(ctx: Context, scratch: torch.Tensor):
    """b200_sort_indices_plan: (took the upper-key-word-first plan, fell back to the 8-pass plan) of the last
    sort_indices call that used `scratch`."""
    hw, fell_back = C.c_int(0), C.c_int(0)
    ctx.check(lib().b200_sort_indices_plan(ctx._h, C.c_void_p(scratch.data_ptr()), C.byref(hw), C.byref(fell_back)))
    return bool(hw.value), bool(fell_back.value)


def sort_keys(ctx: Context, values: Column, order: int = 0) -> torch.Tensor:
    out = torch.empty(max(values.length, 1), dtype=torch.int64, device="cuda:%d" % ctx.device)
    v = values.desc()
    ctx.check(lib().b200_sort_keys(ctx._h, C.byref(v), C.c_int(order), C.c_void_p(out.data_ptr())))
    return out[: values.length]


def sort_keys_rows(ctx: Context, values: Column, order: int = 0, index_base: int = 0):
    """(radix keys, (index_base + row) | class << 62) as two int64-typed tensors."""
    n = values.length
    keys = torch.empty(max(n, 1), dtype=torch.int64, device="cuda:%d" % ctx.device)
    rows = torch.empty(max(n, 1), dtype=torch.int64, device="cuda:%d" % ctx.device)
    v = values.desc()
    ctx.check(lib().b200_sort_keys_rows(ctx._h, C.byref(v), C.c_int(order), C.c_uint64(index_base),
                                        C.c_void_p(keys.data_ptr()), C.c_void_p(rows.data_ptr())))
    return keys[:n], rows[:n]


def sort_indices_chunked(ctx: Context, chunks: Sequence[Column], order: int = 0) -> torch.Tensor:
    """== compute::SortIndices(ChunkedArray, order): indices into the logical concatenation."""
    total = sum(c.length for c in chunks)
    out = torch.empty(max(total, 1), dtype=torch.int64, device="cuda:%d" % ctx.device)
    nb = C.c_size_t(0)
    ctx.check(lib().b200_sort_indices_chunked_scratch_bytes(C.c_int64(total), C.byref(nb)))
    scratch = ctx.empty(nb.value)
    ctx.check(lib().b200_sort_indices_chunked(ctx._h, col_array(chunks), C.c_int(len(chunks)), C.c_int(order),
                                              C.c_void_p(out.data_ptr()), C.c_void_p(scratch.data_ptr()), C.c_size_t(nb.value)))
    return out[:total]


def sort_indices_multi(ctx: Context, keys: Sequence, orders: Sequence[int]) -> torch.Tensor:
    """== compute::SortIndices(RecordBatch | Table, SortOptions{sort_keys}): keys[0] most significant; every key
    puts NaNs then nulls after its values; stable.  A key is a Column, or — every key alike — a sequence of
    Columns (the chunks of a Table column, chunked the same way under every key; indices address the
    concatenation).  int64-typed tensor of uint64 row numbers."""
    chunked = len(keys) > 0 and not isinstance(keys[0], Column)
    ords = (C.c_int * max(len(keys), 1))(*[int(o) for o in orders])
    if chunked:
        n_chunks = len(keys[0])
        if any(len(k) != n_chunks for k in keys):
            raise B200Error(4, "sort_indices_multi: sort keys chunked differently")
        flat = [c for k in keys for c in k]
        n = sum(c.length for c in keys[0])
    else:
        n = keys[0].length if keys else 0
    out = torch.empty(max(n, 1), dtype=torch.int64, device="cuda:%d" % ctx.device)
    nb = C.c_size_t(0)
    if chunked:
        ctx.check(lib().b200_sort_indices_multi_chunked_scratch_bytes(C.c_int64(n), C.byref(nb)))
        scratch = ctx.empty(nb.value)
        ctx.check(lib().b200_sort_indices_multi_chunked(ctx._h, col_array(flat) if flat else None, ords, C.c_int(len(keys)),
                                                        C.c_int(n_chunks), C.c_void_p(out.data_ptr()),
                                                        C.c_void_p(scratch.data_ptr()), C.c_size_t(nb.value)))
        return out[:n]
    ctx.check(lib().b200_sort_indices_multi_scratch_bytes(C.c_int64(n), C.byref(nb)))
    scratch = ctx.empty(nb.value)
    ctx.check(lib().b200_sort_indices_multi(ctx._h, col_array(keys) if keys else None, ords, C.c_int(len(keys)),
                                            C.c_void_p(out.data_ptr()), C.c_void_p(scratch.data_ptr()), C.c_size_t(nb.value)))
    return out[:n]


def sort_range_partition(ctx: Context, values: Column, order: int, index_base: int, splitters: Sequence[int], n_parts: int):
    """b200_sort_range_partition -> (keys, rows, counts): the (radix key, global row | class << 62) pairs of
    the column grouped by destination (stable), int64-typed tensors; `splitters` are unsigned radix keys."""
    n = values.length
    dev = "cuda:%d" % ctx.device
    keys = torch.empty(max(n, 1), dtype=torch.int64, device=dev)
    rows = torch.empty(max(n, 1), dtype=torch.int64, device=dev)
    nb = C.c_size_t(0)
    ctx.check(lib().b200_sort_partition_scratch_bytes(C.c_int64(n), C.byref(nb)))
    scratch = ctx.empty(nb.value)
    spl = (C.c_uint64 * max(n_parts - 1, 1))(*[int(x) & ((1 << 64) - 1) for x in splitters])
    counts = (C.c_int64 * 16)()
    v = values.desc()
    ctx.check(lib().b200_sort_range_partition(ctx._h, C.byref(v), C.c_int(order), C.c_uint64(index_base), spl, C.c_int(n_parts),
                                              C.c_void_p(keys.data_ptr()), C.c_void_p(rows.data_ptr()), counts,
                                              C.c_void_p(scratch.data_ptr()), C.c_size_t(nb.value)))
    return keys[:n], rows[:n], [int(counts[q]) for q in range(n_parts)]


def sort_pairs(ctx: Context, keys: torch.Tensor, payload: torch.Tensor, want_keys: bool = False):
    n = keys.numel()
    nb = C.c_size_t(0)
    ctx.check(lib().b200_sort_pairs_scratch_bytes(C.c_int64(n), C.byref(nb)))
    scratch = ctx.empty(nb.value)
    out_p = torch.empty(max(n, 1), dtype=torch.int64, device=keys.device)
    out_k = torch.empty(max(n, 1), dtype=torch.int64, device=keys.device) if want_keys else None
    ctx.check(lib().b200_sort_pairs(ctx._h, C.c_void_p(keys.data_ptr()), C.c_void_p(payload.data_ptr()), C.c_int64(n),
                                    C.c_void_p(out_k.data_ptr()) if want_keys else None,
                                    C.c_void_p(out_p.data_ptr()), C.c_void_p(scratch.data_ptr()),
                                    C.c_size_t(scratch.numel())))
    return (out_k[:n] if want_keys else None), out_p[:n]


def sort_pairs32(ctx: Context, keys: torch.Tensor, payload32: torch.Tensor, want_keys: bool = False,
                 out: Optional[torch.Tensor] = None):
    """b200_sort_pairs32: stable sort of (uint64 key, uint32 payload) pairs; the payload travels through the
    passes and comes out widened to uint64 (int64-typed tensor; `out`: a caller-owned slice to write it to)."""
    n = keys.numel()
    nb = C.c_size_t(0)
    ctx.check(lib().b200_sort_pairs_scratch_bytes(C.c_int64(n), C.byref(nb)))
    scratch = ctx.empty(nb.value)
    out_p = out if out is not None else torch.empty(max(n, 1), dtype=torch.int64, device=keys.device)
    out_k = torch.empty(max(n, 1), dtype=torch.int64, device=keys.device) if want_keys else None
    ctx.check(lib().b200_sort_pairs32(ctx._h, C.c_void_p(keys.data_ptr()), C.c_void_p(payload32.data_ptr()), C.c_int64(n),
                                      C.c_void_p(out_k.data_ptr()) if want_keys else None,
                                      C.c_void_p(out_p.data_ptr()), C.c_void_p(scratch.data_ptr()),
                                      C.c_size_t(scratch.numel())))
    return (out_k[:n] if want_keys else None), out_p[:n]


class RangePartitionPush:
    """The two phases of the range partition fused with the exchange (b200_sort_range_partition_count / _push)."""

    def __init__(self, ctx: Context, values: Column, order: int, splitters: Sequence[int], n_parts: int):
        self.ctx, self.values, self.order, self.n_parts = ctx, values, order, n_parts
        nb = C.c_size_t(0)
        ctx.check(lib().b200_sort_partition_scratch_bytes(C.c_int64(values.length), C.byref(nb)))
        self.scratch = ctx.empty(nb.value)
        self.spl = (C.c_uint64 * max(n_parts - 1, 1))(*[int(x) & ((1 << 64) - 1) for x in splitters])
        counts = (C.c_int64 * 16)()
        v = values.desc()
        ctx.check(lib().b200_sort_range_partition_count(ctx._h, C.byref(v), C.c_int(order), self.spl, C.c_int(n_parts), counts,
                                                        C.c_void_p(self.scratch.data_ptr()), C.c_size_t(self.scratch.numel())))
        self.counts = [int(counts[q]) for q in range(n_parts)]

    def push(self, index_base: int, dst_keys: Sequence[int], dst_rows: Sequence[int], dst_offsets: Sequence[int], rows32: bool):
        """dst_keys / dst_rows: device addresses (possibly peer mappings), one per destination."""
        n = self.n_parts
        kp = (C.c_void_p * n)(*[int(x) for x in dst_keys])
        rp = (C.c_void_p * n)(*[int(x) for x in dst_rows])
        off = (C.c_int64 * n)(*[int(x) for x in dst_offsets])
        v = self.values.desc()
        self.ctx.check(lib().b200_sort_range_partition_push(
            self.ctx._h, C.byref(v), C.c_int(self.order), C.c_uint64(index_base), self.spl, C.c_int(n), kp, rp, off,
            C.c_int(1 if rows32 else 0), C.c_void_p(self.scratch.data_ptr()), C.c_size_t(self.scratch.numel())))


class _RawCuda:
    """Borrowed device memory as a __cuda_array_interface__ provider (zero-copy torch view)."""

    def __init__(self, ptr: int, nbytes: int, owner):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}
        self._owner = owner


class PeerBuffer:
    """A device allocation other processes of the node can map (b200_peer_alloc / _export / _open)."""

    def __init__(self, ctx: Context, nbytes: int):
        self.ctx, self.nbytes = ctx, int(nbytes)
        p = C.c_void_p()
        ctx.check(lib().b200_peer_alloc(ctx._h, C.c_size_t(self.nbytes), C.byref(p)))
        self.ptr = int(p.value)
        self.peers: List[int] = []     # mappings of the other ranks' buffers opened here

    def handle(self) -> bytes:
        h = (C.c_ubyte * 64)()
        self.ctx.check(lib().b200_peer_export(self.ctx._h, C.c_void_p(self.ptr), h))
        return bytes(h)

    def open(self, handle: bytes) -> int:
        h = (C.c_ubyte * 64)(*handle)
        p = C.c_void_p()
        self.ctx.check(lib().b200_peer_open(self.ctx._h, h, C.byref(p)))
        self.peers.append(int(p.value))
        return int(p.value)

    def tensor(self, dtype: torch.dtype, count: int, byte_offset: int = 0) -> torch.Tensor:
        item = torch.empty(0, dtype=dtype).element_size()
        assert byte_offset + count * item <= self.nbytes
        raw = torch.as_tensor(_RawCuda(self.ptr + byte_offset, max(count * item, 1), self), device="cuda:%d" % self.ctx.device)
        return raw[: count * item].view(dtype)

    def close_peers(self):
        for p in self.peers:
            lib().b200_peer_close(self.ctx._h, C.c_void_p(p))
        self.peers = []

    def close(self):
        self.close_peers()
        if self.ptr:
            lib().b200_peer_free(self.ctx._h, C.c_void_p(self.ptr))
            self.ptr = 0


# ---------------------------------------------------------------------------------------------
# Grouper + hash aggregates


class GroupByOutput:
    """aggs[j] = list of Columns (one, or [mins, maxes] for hash_min_max); keys[k] = Column."""

    def __init__(self, num_groups: int, aggs: List[List[Column]], keys: List[Column]):
        self.num_groups, self.aggs, self.keys = num_groups, aggs, keys


class GroupBy:
    """== the streaming state of compute::internal::GroupBy (hash_aggregate.cc:890-966)."""

    def __init__(self, ctx: Context, key_types: Sequence[int], aggs: Sequence[AggSpec], arg_types: Sequence[int],
                 expected_groups: int = 0):
        self.ctx = ctx
        self.key_types = list(key_types)
        self.aggs = list(aggs)
        self.arg_types = list(arg_types)
        kt = (C.c_int32 * max(len(key_types), 1))(*key_types)
        specs = (b200_agg_spec * max(len(aggs), 1))()
        for j, a in enumerate(aggs):
            specs[j] = b200_agg_spec(a.kind, arg_types[j], a.option, 0)
        self._h = C.c_void_p()
        ctx.check(lib().b200_groupby_create(ctx._h, C.c_int(len(key_types)), kt, C.c_int(len(aggs)), specs,
                                            C.c_int64(expected_groups), C.byref(self._h)))

    def close(self):
        if self._h:
            lib().b200_groupby_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_option(self, name: str, value: int):
        self.ctx.check(lib().b200_groupby_set_option(self._h, name.encode(), C.c_int64(value)))

    def stat(self, name: str) -> int:
        v = C.c_int64(0)
        self.ctx.check(lib().b200_groupby_stat(self._h, name.encode(), C.byref(v)))
        return int(v.value)

    def consume(self, keys: Sequence[Column], args: Sequence[Column], row_base: int = 0):
        self.ctx.check(lib().b200_groupby_consume(self._h, col_array(keys), col_array(args), C.c_uint64(row_base)))

    def consume_selected(self, keys: Sequence[Column], args: Sequence[Column], selection: Column, row_base: int = 0):
        """b200_groupby_consume_selected: only the rows whose bit is set in `selection` (a boolean device Column
        at offset 0 without nulls, e.g. compare_scalar's output on a column without nulls)."""
        if selection.type != BOOL or selection.offset != 0 or (selection.validity is not None and selection.null_count != 0):
            raise B200Error(10, "consume_selected: the selection must be a boolean column at offset 0 without nulls")
        if selection.length != keys[0].length:
            raise B200Error(4, "Array arguments must all be the same length")
        self.ctx.check(lib().b200_groupby_consume_selected(self._h, col_array(keys), col_array(args), C.c_uint64(row_base),
                                                           C.c_void_p(selection.data.data_ptr())))

    def consume_fused(self, keys: Sequence[Column], args: Sequence, predicate=None, row_base: int = 0):
        """b200_groupby_consume_fused.  args[j] is a Column, or a tuple (op1, a, op2, scalar, b) meaning
        `a op1 (scalar op2 b)` over float64 columns; predicate = (column, op name, scalar) or None."""
        from .columns import b200_fused_arg, b200_predicate
        arr = (b200_fused_arg * max(len(args), 1))()
        for j, g in enumerate(args):
            if isinstance(g, Column):
                arr[j].kind = 0
                arr[j].a = g.desc()
            else:
                op1, a, op2, scalar, b = g
                arr[j].kind, arr[j].op1, arr[j].op2 = 1, ARITH_OPS[op1], ARITH_OPS[op2]
                arr[j].a, arr[j].b, arr[j].scalar = a.desc(), b.desc(), float(scalar)
        pred, keep = None, None
        if predicate is not None:
            col, op, scalar = predicate
            keep = np.array([scalar], dtype=NP_OF[col.type])
            pred = b200_predicate(col.desc(), CMP_OPS[op], 0, keep.ctypes.data)
        self.ctx.check(lib().b200_groupby_consume_fused(self._h, col_array(keys), arr, C.c_uint64(row_base),
                                                        C.byref(pred) if pred is not None else None))

    def consume_rows(self, keys: Sequence[Column], args: Sequence[Column], row_ids: torch.Tensor):
        """b200_groupby_consume_rows: explicit global row numbers (int64-typed tensor of uint64)."""
        self.ctx.check(lib().b200_groupby_consume_rows(self._h, col_array(keys), col_array(args),
                                                       C.c_void_p(row_ids.data_ptr())))

    def first_rows(self) -> torch.Tensor:
        """Global first-occurrence row of every group, in group order (int64-typed uint64)."""
        g = self.num_groups()
        out = torch.empty(max(g, 1), dtype=torch.int64, device="cuda:%d" % self.ctx.device)
        self.ctx.check(lib().b200_groupby_first_rows(self._h, C.c_void_p(out.data_ptr())))
        return out[:g]

    def partition_plan(self, keys: Sequence[Column], n_parts: int):
        """b200_groupby_partition_plan -> (pos uint32-as-int32 tensor [n], counts list[n_parts])."""
        n = keys[0].length
        dev = "cuda:%d" % self.ctx.device
        nbytes = C.c_size_t(0)
        self.ctx.check(lib().b200_partition_scratch_bytes(C.c_int64(n), C.byref(nbytes)))
        scratch = self.ctx.empty(nbytes.value)
        pos = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
        counts = (C.c_int64 * 16)()
        self.ctx.check(lib().b200_groupby_partition_plan(self._h, col_array(keys), C.c_int(n_parts), C.c_void_p(pos.data_ptr()),
                                                         counts, C.c_void_p(scratch.data_ptr()), C.c_size_t(nbytes.value)))
        return pos[:n], [int(counts[p]) for p in range(n_parts)]

    def grouper_consume(self, keys: Sequence[Column]) -> torch.Tensor:
        n = keys[0].length
        ids = torch.empty(max(n, 1), dtype=torch.int32, device="cuda:%d" % self.ctx.device)
        self.ctx.check(lib().b200_grouper_consume(self._h, col_array(keys), C.c_void_p(ids.data_ptr())))
        return ids[:n]

    def agg_consume(self, agg_index: int, arg: Column, ids: torch.Tensor, num_groups: int):
        a = arg.desc()
        self.ctx.check(lib().b200_hash_agg_consume(self._h, C.c_int(agg_index), C.byref(a),
                                                   C.c_void_p(ids.data_ptr()), C.c_uint32(num_groups)))

    def num_groups(self) -> int:
        g = C.c_uint32(0)
        self.ctx.check(lib().b200_groupby_num_groups(self._h, C.byref(g)))
        return int(g.value)

    def merge(self, other: "GroupBy"):
        self.ctx.check(lib().b200_groupby_merge(self._h, other._h))

    def slot_bytes(self) -> int:
        n = C.c_size_t(0)
        self.ctx.check(lib().b200_groupby_state_layout(self._h, C.byref(n)))
        return int(n.value)

    def export_state(self) -> torch.Tensor:
        g = self.num_groups()
        sb = self.slot_bytes()
        buf = self.ctx.empty(max(g, 1) * sb)
        ng = C.c_uint32(0)
        self.ctx.check(lib().b200_groupby_export(self._h, C.c_void_p(buf.data_ptr()), C.byref(ng)))
        return buf[: ng.value * sb]

    def import_state(self, state: torch.Tensor):
        sb = self.slot_bytes()
        n = state.numel() // sb
        self.ctx.check(lib().b200_groupby_import(self._h, C.c_void_p(state.data_ptr()), C.c_uint32(n)))

    def finalize(self) -> GroupByOutput:
        ctx = self.ctx
        G = self.num_groups()
        out_types = [agg_out_types(a.kind, t) for a, t in zip(self.aggs, self.arg_types)]
        flat = [t for ts in out_types for t in ts]
        out_aggs = (b200_out_col * max(len(flat), 1))()
        bufs = []
        for i, t in enumerate(flat):
            bits = BITS[t]
            d = ctx.empty(max(G, 1) * max(bits // 8, 1))
            v = ctx.empty((G + 7) // 8 + 8)
            out_aggs[i] = b200_out_col(v.data_ptr(), d.data_ptr(), t, 0, 0, 0)
            bufs.append((t, d, v))
        out_keys = (b200_out_col * max(len(self.key_types), 1))()
        kbufs = []
        for i, t in enumerate(self.key_types):
            x = None
            if t == BINARY:   # int32 offsets + the value bytes, sized by the library
                nbytes = C.c_int64(0)
                ctx.check(lib().b200_groupby_key_bytes(self._h, C.c_int(i), C.byref(nbytes)))
                d = ctx.empty((G + 1) * 4 + 8)
                x = ctx.empty(max(int(nbytes.value), 1) + 8)
            else:
                bits = BITS[t]
                d = ctx.empty((G + 7) // 8 + 8 if bits == 1 else max(G, 1) * (bits // 8))
            v = ctx.empty((G + 7) // 8 + 8)
            out_keys[i] = b200_out_col(v.data_ptr(), d.data_ptr(), t, 0, 0, 0, x.data_ptr() if x is not None else None)
            kbufs.append((t, d, v, x))
        ctx.check(lib().b200_groupby_finalize(self._h, out_aggs, out_keys))
        ctx.synchronize()
        aggs, i = [], 0
        for ts in out_types:
            cols = []
            for _ in ts:
                t, d, v = bufs[i]
                cols.append(Column(t, d, v, G, 0, int(out_aggs[i].null_count)))
                i += 1
            if len(cols) == 2:
                cols[1] = Column(cols[1].type, cols[1].data, cols[0].validity, G, 0, cols[0].null_count)
            aggs.append(cols)
        keys = [Column(t, d, v, G, 0, int(out_keys[k].null_count), x) for k, (t, d, v, x) in enumerate(kbufs)]
        return GroupByOutput(G, aggs, keys)


def partition_scatter(ctx: Context, col: Column, pos: torch.Tensor, want_valid: bool):
    """b200_partition_scatter -> (data bytes tensor, validity bytes tensor or None), partitioned layout."""
    n = col.length
    width = BITS[col.type] // 8
    out = ctx.empty(max(n, 1) * width)
    valid = ctx.empty(max(n, 1)) if want_valid else None
    d = col.desc()
    ctx.check(lib().b200_partition_scatter(ctx._h, C.byref(d), C.c_void_p(pos.data_ptr()), C.c_void_p(out.data_ptr()),
                                           C.c_void_p(valid.data_ptr()) if want_valid else None))
    return out[: n * width], (valid[:n] if want_valid else None)


def partition_rows(ctx: Context, n: int, row_base: int, pos: torch.Tensor) -> torch.Tensor:
    out = torch.empty(max(n, 1), dtype=torch.int64, device="cuda:%d" % ctx.device)
    ctx.check(lib().b200_partition_rows(ctx._h, C.c_int64(n), C.c_uint64(row_base), C.c_void_p(pos.data_ptr()),
                                        C.c_void_p(out.data_ptr())))
    return out[:n]


def pack_validity(ctx: Context, valid_bytes: torch.Tensor):
    """b200_pack_validity -> (bitmap bytes tensor, null_count)."""
    n = valid_bytes.numel()
    bitmap = ctx.empty((n + 31) // 32 * 4 + 8)
    nulls = C.c_int64(0)
    ctx.check(lib().b200_pack_validity(ctx._h, C.c_void_p(valid_bytes.data_ptr()), C.c_int64(n), C.c_void_p(bitmap.data_ptr()),
                                       C.byref(nulls)))
    return bitmap, int(nulls.value)


def make_groupings(ctx: Context, ids: torch.Tensor, num_groups: int):
    """== Grouper::MakeGroupings (hash_aggregate.cc:968-1001): ids = uint32 group ids as an int32-typed device
    tensor.  Returns (offsets int32[num_groups + 1], indices int32[n]) — the buffers of a list<int32> array."""
    n = ids.numel()
    dev = "cuda:%d" % ctx.device
    offsets = torch.empty(num_groups + 1, dtype=torch.int32, device=dev)
    indices = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
    nbytes = C.c_size_t(0)
    ctx.check(lib().b200_make_groupings_scratch_bytes(C.c_int64(n), C.byref(nbytes)))
    scratch = ctx.empty(nbytes.value)
    ctx.check(lib().b200_make_groupings(ctx._h, C.c_void_p(ids.data_ptr()), C.c_int64(n), C.c_uint32(num_groups),
                                        C.c_void_p(offsets.data_ptr()), C.c_void_p(indices.data_ptr()),
                                        C.c_void_p(scratch.data_ptr()), C.c_size_t(scratch.numel())))
    return offsets, indices[:n]


def dictionary_indices(ctx: Context, ids: torch.Tensor, null_group: int = -1):
    """== DictEncodeAction (vector_hash.cc:165-226): uint32 group ids (int32-typed device tensor) -> (int32 indices,
    validity bitmap bytes or None).  null_group >= 0: DictionaryEncodeOptions::MASK with a null in the column."""
    n = ids.numel()
    dev = "cuda:%d" % ctx.device
    out = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
    validity = torch.zeros(((n + 31) // 32) * 4 + 64, dtype=torch.uint8, device=dev) if null_group >= 0 else None
    ctx.check(lib().b200_dictionary_indices(ctx._h, C.c_void_p(ids.data_ptr()), C.c_int64(n), C.c_int64(null_group),
                                            C.c_void_p(out.data_ptr()),
                                            C.c_void_p(validity.data_ptr()) if validity is not None else None))
    return out[:n], validity


def group_by(ctx: Context, keys: Sequence[Column], aggs: Sequence[AggSpec], args: Sequence[Column],
             expected_groups: int = 0, row_base: int = 0, selection: Optional[Column] = None,
             options: Optional[dict] = None) -> GroupByOutput:
    """== compute::internal::GroupBy(arguments, keys, aggregates) in one batch; with `selection`, of the selected
    rows only (== GroupBy over Filter(column, selection) of every column)."""
    gb = GroupBy(ctx, [k.type for k in keys], aggs, [a.type for a in args], expected_groups)
    try:
        for name, value in (options or {}).items():
            gb.set_option(name, value)
        if selection is None:
            gb.consume(keys, args, row_base)
        else:
            gb.consume_selected(keys, args, selection, row_base)
        return gb.finalize()
    finally:
        gb.close()
