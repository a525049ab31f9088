"""Host-side mirror of the reference's public entry points for the hot path.

    Filter(values, mask, null_selection)      == compute::Filter        (compute/api_vector.cc:94-99)
    Take(values, indices, boundscheck)        == compute::Take          (compute/api_vector.cc:101-105)
    SortIndices(values, order)                == compute::SortIndices   (compute/api_vector.cc:48-53)
    GroupBy(arguments, keys, aggregates)      == compute::internal::GroupBy
                                                 (compute/kernels/hash_aggregate.cc:890-966)

Same argument meaning and error behaviour (status codes of arrow::StatusCode carried by B200Error).
Inputs are `Column`s; when they live in host memory (numpy buffers) they are staged to the GPU,
the C-ABI kernels run there, and the result comes back to host memory: that round trip is what
bench.py's `e2e` leg times.  Device-resident Columns (torch buffers, the analogue of
arrow_cuda CudaBuffer-backed ArrayData) stay on the device.  Nothing here computes on the CPU.
"""
from __future__ import annotations

import weakref
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import capi
from .capi import B200Error
from .columns import BITS, AggSpec, Column

DROP, EMIT_NULL = 0, 1          # FilterOptions::NullSelectionBehavior (api_vector.h:37-49)
ASCENDING, DESCENDING = 0, 1    # SortOrder (api_vector.h:61-66)

_contexts: Dict[int, capi.Context] = {}


def default_context(device: Optional[int] = None) -> capi.Context:
    """One shared context per device, the role default_exec_context() plays (compute/exec.cc:61-64)."""
    if device is None:
        device = torch.cuda.current_device() if torch.cuda.is_available() else 0
    ctx = _contexts.get(device)
    if ctx is None:
        ctx = capi.Context(device)
        _contexts[device] = ctx
    return ctx


class PinnedPool:
    """Page-locked host staging buffers, reused across calls (cudaHostAlloc is slow)."""

    def __init__(self):
        self._free: Dict[int, List[torch.Tensor]] = {}

    def get(self, nbytes: int) -> torch.Tensor:
        size = 1 << max(int(nbytes - 1).bit_length(), 12)
        lst = self._free.get(size)
        if lst:
            return lst.pop()
        return torch.empty(size, dtype=torch.uint8, pin_memory=True)

    def put(self, t: torch.Tensor) -> None:
        self._free.setdefault(t.numel(), []).append(t)


_pool = PinnedPool()


def pinned_column(col: Column) -> Column:
    """Copy of a host column whose buffers are page-locked (so H2D copies run at PCIe speed)."""
    assert not col.on_device
    d = _pool.get(col.data.nbytes)
    d[: col.data.nbytes].numpy()[:] = col.data
    v = None
    if col.validity is not None:
        v = _pool.get(col.validity.nbytes)
        v[: col.validity.nbytes].numpy()[:] = col.validity
    return Column(col.type, d[: col.data.nbytes].numpy(), None if v is None else v[: col.validity.nbytes].numpy(),
                  col.length, col.offset, col.null_count)


def _from_host(col: Column) -> bool:
    """True for host columns and for columns Pipeline.stage() copied to the device ahead of their call."""
    return (not col.on_device) or getattr(col, "from_host", False)


def _h2d(col: Column, ctx: capi.Context) -> Tuple[Column, int]:
    if col.on_device:
        return col, getattr(col, "staged_bytes", 0)
    dev = "cuda:%d" % ctx.device
    d = torch.from_numpy(col.data).to(dev, non_blocking=True)
    v = None
    n = col.data.nbytes
    if col.validity is not None:
        v = torch.from_numpy(col.validity).to(dev, non_blocking=True)
        n += col.validity.nbytes
    x = None
    if col.extra is not None:   # BINARY key column: the value bytes
        x = torch.from_numpy(col.extra).to(dev, non_blocking=True)
        n += col.extra.nbytes
    return Column(col.type, d, v, col.length, col.offset, col.null_count, x), n


def _d2h(t: torch.Tensor, nbytes: int) -> np.ndarray:
    """Device bytes -> page-locked host buffer (returned as a numpy view; the caller owns it)."""
    h = _pool.get(max(nbytes, 1))
    h[:nbytes].copy_(t.view(torch.uint8).reshape(-1)[:nbytes], non_blocking=True)
    arr = h[:nbytes].numpy()
    weakref.finalize(arr, _pool.put, h)  # the staging buffer is recycled when the result dies
    return arr


class Transfer:
    """Byte counts of the last call's staging copies (bench.py's h2d/d2h_bytes_per_step)."""
    h2d = 0
    d2h = 0


def Filter(values: Column, mask: Column, null_selection: int = DROP, ctx: Optional[capi.Context] = None,
           _wait: bool = True) -> Column:
    ctx = ctx or default_context()
    host = _from_host(values)
    dv, b0 = _h2d(values, ctx)
    dm, b1 = _h2d(mask, ctx)
    out = capi.filter(ctx, dv, dm, null_selection)
    Transfer.h2d, Transfer.d2h = b0 + b1, 0
    if not host:
        return out.col
    col = out.col
    bits = BITS[col.type]
    nd = (col.length + 7) // 8 if bits == 1 else col.length * (bits // 8)
    data = _d2h(col.data, nd)
    validity = None
    nv = 0
    if out.has_validity:
        nv = (col.length + 7) // 8
        validity = _d2h(col.validity, nv)
    if _wait:
        ctx.stream.synchronize()
    Transfer.d2h = nd + nv + 8
    return Column(col.type, data, validity, col.length, 0, col.null_count)


def SortIndices(values, order=ASCENDING, ctx: Optional[capi.Context] = None, _wait: bool = True):
    """== compute::SortIndices.  `values` is a Column (order: ASCENDING / DESCENDING), or a sequence of Columns
    with a sequence of orders — the sort keys of SortIndices(RecordBatch | Table, SortOptions), most significant
    first (vector_sort.cc:1632-1711).  uint64 indices: numpy array for host input, int64-typed torch tensor for
    device input."""
    if not isinstance(values, Column):   # argument errors come before any device work
        keys = list(values)
        if len(keys) == 0:
            raise B200Error(4, "Must specify one or more sort keys")
        orders = [order] * len(keys) if isinstance(order, int) else list(order)
        if len(orders) != len(keys):
            raise B200Error(4, "SortIndices: one order per sort key")
    ctx = ctx or default_context()
    if isinstance(values, Column):
        host = _from_host(values)
        dv, b0 = _h2d(values, ctx)
        out = capi.sort_indices(ctx, dv, order)
        n = values.length
    else:
        host = _from_host(keys[0])
        staged = [_h2d(k, ctx) for k in keys]
        b0 = sum(b for _, b in staged)
        out = capi.sort_indices_multi(ctx, [c for c, _ in staged], orders)
        n = keys[0].length
    Transfer.h2d, Transfer.d2h = b0, 0
    if not host:
        return out
    h = _d2h(out, n * 8)
    if _wait:
        ctx.stream.synchronize()
    Transfer.d2h = n * 8
    return h.view(np.uint64)


def Take(values, indices, boundscheck: bool = True, ctx: Optional[capi.Context] = None) -> Column:
    """== compute::Take (compute/api_vector.cc:101-105) for fixed-width values.  `values` is a Column or a
    sequence of Columns (ChunkedArray values, TakeCA vector_selection.cc:1921-1944: indices address the
    concatenation; the result is one Column).  Host columns are staged to the device and the result comes back
    to the host; device columns (or a device index tensor from SortIndices) stay."""
    if not isinstance(values, Column) and len(values) == 0:
        raise B200Error(4, "Must pass at least one array")   # Concatenate's message on TakeCA's route
    ctx = ctx or default_context()
    if isinstance(values, Column):
        host = not values.on_device
        dv, b0 = _h2d(values, ctx)
    else:
        host = not values[0].on_device
        staged = [_h2d(v, ctx) for v in values]
        dv, b0 = [c for c, _ in staged], sum(b for _, b in staged)
    di, b1 = (indices, 0) if isinstance(indices, torch.Tensor) else _h2d(indices, ctx)
    out = capi.take(ctx, dv, di, boundscheck)
    Transfer.h2d, Transfer.d2h = b0 + b1, 0
    if not host:
        return out
    bits = BITS[out.type]
    nd = (out.length + 7) // 8 if bits == 1 else out.length * (bits // 8)
    nv = (out.length + 7) // 8
    data, validity = _d2h(out.data, nd), _d2h(out.validity, nv)
    ctx.stream.synchronize()
    Transfer.d2h = nd + nv
    return Column(out.type, data, validity, out.length, 0, out.null_count)


class Pending:
    """Result of Pipeline.submit: `result()` blocks until the lane's copies and kernels are done."""

    def __init__(self, event: torch.cuda.Event, value):
        self._event, self._value = event, value

    def done(self) -> bool:
        return self._event.query()

    def result(self):
        self._event.synchronize()
        return self._value


class Pipeline:
    """Several calls in flight on one GPU: `depth` lanes, each one b200_ctx on its own CUDA stream
    ("callers wanting concurrency create one b200_ctx per stream").  A call submitted
    to a lane stages its host inputs, runs and reads its result back on that lane's stream only, so the
    H2D copy of call i+1, the kernels of call i and the D2H copy of call i-1 overlap: both PCIe
    directions and the SMs are busy at once, which is what bounds an ingest loop over host batches.
    The reference's counterpart is calling compute::Filter / SortIndices from several CPU threads
    (its kernels run on the calling thread, compute/exec.cc:972-986).

        pipe = Pipeline(depth=3)
        pending = [pipe.submit(SortIndices, batch, ASCENDING) for batch in batches]
        for p in pending: indices = p.result()
    """

    def __init__(self, depth: int = 3, device: Optional[int] = None):
        if device is None:
            device = torch.cuda.current_device()
        self.device = device
        self.lanes = []
        for _ in range(max(depth, 1)):
            stream = torch.cuda.Stream(device)
            self.lanes.append((capi.Context(device, stream), stream))
        self._next = 0
        self._last: List[Optional[torch.cuda.Event]] = [None] * len(self.lanes)

    @property
    def launches(self) -> int:
        return sum(ctx.launches for ctx, _ in self.lanes)

    def stage(self, *cols: Column):
        """Starts the host -> device copies of one call's host columns on the next lane and returns at once; pass
        the results to submit() in place of the host columns.  Lets a caller start the copies of call i+1 before
        call i reaches a point where it has to wait for the device (an output length, a collective's counts).
        Returns one staged column, or a tuple of them."""
        i = self._next % len(self.lanes)
        self._next += 1
        ctx, stream = self.lanes[i]
        out = []
        with torch.cuda.stream(stream):
            for col in cols:
                dev, nbytes = _h2d(col, ctx)
                dev.from_host, dev.staged_bytes, dev.lane = True, nbytes, i
                out.append(dev)
        return out[0] if len(out) == 1 else tuple(out)

    def submit(self, fn, *args, **kwargs) -> Pending:
        """fn is Filter or SortIndices (calls that need a host-visible size mid-way, GroupBy's group count
        or Filter's output length, block the submitting thread for that long but still overlap with the
        other lanes' copies).  A staged column among the arguments pins the call to the lane that holds it."""
        staged = [a for a in args if isinstance(a, Column) and getattr(a, "lane", None) is not None]
        if staged:
            i = staged[0].lane
        else:
            i = self._next % len(self.lanes)
            self._next += 1
        ctx, stream = self.lanes[i]
        with torch.cuda.stream(stream):
            if fn in (Filter, SortIndices) or getattr(fn, "_b200_async", False):
                kwargs["_wait"] = False
            value = fn(*args, ctx=ctx, **kwargs)
            ev = torch.cuda.Event()
            ev.record(stream)
        self._last[i] = ev
        return Pending(ev, value)

    def join(self, stream: Optional[torch.cuda.Stream] = None) -> None:
        """Make `stream` (default: the current one) wait for everything submitted so far."""
        stream = stream or torch.cuda.current_stream(self.device)
        for ev in self._last:
            if ev is not None:
                stream.wait_event(ev)

    def close(self):
        for ctx, stream in self.lanes:
            stream.synchronize()
            ctx.close()
        self.lanes = []


DOUBLE_T = 12   # columns.DOUBLE
Q1_SHIPDATE_CUTOFF = 10471   # 1998-09-02 as date32 (TPC-H Q1: l_shipdate <= date '1998-12-01' - 90 days)


def q1_aggregates():
    """The Q1 outputs: argument index into (quantity, extendedprice, discount,
    disc_price = extendedprice * (1 - discount)) and the aggregate applied to it."""
    from . import columns as cols
    return [(0, AggSpec(cols.HASH_SUM)), (1, AggSpec(cols.HASH_SUM)), (3, AggSpec(cols.HASH_SUM)),
            (0, AggSpec(cols.HASH_MEAN)), (1, AggSpec(cols.HASH_MEAN)), (2, AggSpec(cols.HASH_MEAN)),
            (0, AggSpec(cols.HASH_COUNT))]


def Q1Pipeline(shipdate: Column, returnflag: Column, linestatus: Column, measures: Sequence[Column],
               ctx: Optional[capi.Context] = None):
    """TPC-H Q1-shaped pipeline through the public operations, unfused (config C5):
        mask = less_equal(shipdate, cutoff)                   scalar_compare.cc
        c'   = Filter(c, mask) for every used column          one plan shared by all columns
                                                              (FilterRecordBatch, vector_selection.cc:1776-1866)
        disc_price' = multiply(extendedprice', subtract(1.0, discount'))      scalar_arithmetic.cc
        GroupBy(measures' + [disc_price'], keys = (returnflag', linestatus'), q1_aggregates())
    Device columns in, device GroupByOutput out."""
    ctx = ctx or default_context()
    mask = capi.compare_scalar(ctx, shipdate, "less_equal", Q1_SHIPDATE_CUTOFF)
    plan = capi.FilterPlan(ctx, mask, DROP, sync=True)
    rf, ls = plan.apply(returnflag).col, plan.apply(linestatus).col
    ms = [plan.apply(m).col for m in measures]
    ms.append(capi.arithmetic(ctx, "multiply", ms[1], capi.arithmetic(ctx, "subtract", 1.0, ms[2])))
    specs = q1_aggregates()
    return capi.group_by(ctx, [rf, ls], [a for _, a in specs], [ms[i] for i, _ in specs], expected_groups=8)


def Q1PipelineSelected(shipdate: Column, returnflag: Column, linestatus: Column, measures: Sequence[Column],
                       ctx: Optional[capi.Context] = None):
    """The same query with the Filter folded into the aggregation: the comparison's bitmap is
    the selection of b200_groupby_consume_selected, so no filtered column is materialised — 38 B/row of traffic
    instead of ~100.  Same result as Q1Pipeline (group order included); shipdate must not hold nulls."""
    ctx = ctx or default_context()
    if shipdate.validity is not None and shipdate.null_count != 0:
        return Q1Pipeline(shipdate, returnflag, linestatus, measures, ctx)
    mask = capi.compare_scalar(ctx, shipdate, "less_equal", Q1_SHIPDATE_CUTOFF)
    ms = list(measures)
    ms.append(capi.arithmetic(ctx, "multiply", ms[1], capi.arithmetic(ctx, "subtract", 1.0, ms[2])))
    specs = q1_aggregates()
    return capi.group_by(ctx, [returnflag, linestatus], [a for _, a in specs], [ms[i] for i, _ in specs], expected_groups=8,
                         selection=mask)


def Q1PipelineFused(shipdate: Column, returnflag: Column, linestatus: Column, measures: Sequence[Column],
                    ctx: Optional[capi.Context] = None):
    """The same query in ONE pass over the inputs (b200_groupby_consume_fused): the predicate, the derived measure
    extendedprice * (1 - discount) and all seven aggregates are evaluated while shipdate, the two key columns and
    the three measures stream through once (30 B/row).  Same result as Q1Pipeline, group order included."""
    ctx = ctx or default_context()
    qty, price, disc = measures
    derived = ("multiply", price, "subtract", 1.0, disc)
    stream = [qty, price, disc, derived]
    specs = q1_aggregates()
    aggs = [a for _, a in specs]
    gb = capi.GroupBy(ctx, [returnflag.type, linestatus.type], aggs, [DOUBLE_T] * len(aggs), 8)
    try:
        gb.consume_fused([returnflag, linestatus], [stream[i] for i, _ in specs], (shipdate, "less_equal", Q1_SHIPDATE_CUTOFF))
        return gb.finalize()
    finally:
        gb.close()


class GroupByResult:
    """Image of GroupBy's StructArray (hash_aggregate.cc:962-965): one entry per aggregate (a list
    of (values, valid) pairs: two for hash_min_max) then the key columns, `num_groups` long."""

    def __init__(self, num_groups, aggs, keys):
        self.num_groups, self.aggs, self.keys = num_groups, aggs, keys


def GroupBy(arguments: Sequence[Column], keys: Sequence[Column], aggregates: Sequence[AggSpec],
            ctx: Optional[capi.Context] = None, expected_groups: int = 0):
    ctx = ctx or default_context()
    host = not keys[0].on_device
    nbytes = 0
    dkeys, dargs, seen = [], [], {}
    for c in list(keys) + list(arguments):
        key = (id(c.data), id(c.validity), c.offset)
        if key not in seen:
            seen[key], b = _h2d(c, ctx)
            nbytes += b
    dkeys = [seen[(id(c.data), id(c.validity), c.offset)] for c in keys]
    dargs = [seen[(id(c.data), id(c.validity), c.offset)] for c in arguments]
    out = capi.group_by(ctx, dkeys, aggregates, dargs, expected_groups=expected_groups)
    Transfer.h2d, Transfer.d2h = nbytes, 0
    if not host:
        return out
    d2h = 0

    def pair(c):
        """device result column -> (values, valid) on the host through the pinned pool (DMA, no pageable staging)"""
        nonlocal d2h
        from .columns import BINARY, BITS, NP_OF, BOOL, unpack_bits
        g = c.length
        if c.type == BINARY:   # a variable-length key column comes back as a host Column (offsets + bytes)
            h = c.to_host()
            d2h += (g + 1) * 4 + (int(h.data[: (g + 1) * 4].view(np.int32)[g]) if g else 0) + (g + 7) // 8
            return c.type, h, None, g
        bits = BITS[c.type]
        nb = (g + 7) // 8 if bits == 1 else g * (bits // 8)
        data = _d2h(c.data, nb)
        vbytes = _d2h(c.validity, (g + 7) // 8) if c.validity is not None else None
        d2h += nb + (0 if vbytes is None else (g + 7) // 8)
        return c.type, data, vbytes, g

    staged_aggs = [[pair(c) for c in cols_] for cols_ in out.aggs]
    staged_keys = [pair(c) for c in out.keys]
    ctx.stream.synchronize()

    def finish(t, data, vbytes, g):
        from .columns import BINARY, BITS, NP_OF, BOOL, unpack_bits
        if t == BINARY:
            vals = np.empty(g, dtype=object)
            for i, x in enumerate(data.strings()):
                vals[i] = b"" if x is None else x
            return vals, data.valid()
        vals = unpack_bits(data, g) if t == BOOL else data[: g * (BITS[t] // 8)].view(NP_OF[t])
        valid = np.ones(g, dtype=np.bool_) if vbytes is None else unpack_bits(vbytes, g)
        return vals, valid

    res_aggs = [[finish(*p) for p in lst] for lst in staged_aggs]
    res_keys = [finish(*p) for p in staged_keys]
    Transfer.d2h = d2h
    return GroupByResult(out.num_groups, res_aggs, res_keys)


# ---- other consumers of the Grouper -----------------------------------------------------------
def MakeGroupings(ids, num_groups: int, ctx: Optional[capi.Context] = None):
    """== Grouper::MakeGroupings (hash_aggregate.cc:968-1001): uint32 group ids (device tensor) -> the offsets and
    values buffers of the list<int32> array whose slot g lists the rows holding id g."""
    ctx = ctx or default_context()
    return capi.make_groupings(ctx, ids, num_groups)


def ApplyGroupings(groupings, array: Column, ctx: Optional[capi.Context] = None):
    """== Grouper::ApplyGroupings (hash_aggregate.cc:1003-1018): Take(array, groupings.values) under the same offsets.
    Returns (offsets, values Column)."""
    ctx = ctx or default_context()
    offsets, indices = groupings
    from .columns import INT32
    idx = Column(INT32, indices.view(torch.uint8), None, indices.numel(), 0, 0)
    return offsets, capi.take(ctx, array, idx, True)


def Unique(values: Column, ctx: Optional[capi.Context] = None) -> Column:
    """== "unique" (compute/kernels/vector_hash.cc:723-782) for fixed-width columns: the distinct values in
    first-occurrence order, a null (if any) as one null entry — the Grouper's GetUniques over the one column."""
    ctx = ctx or default_context()
    out = capi.group_by(ctx, [values], [], [])
    return out.keys[0]


def ValueCounts(values: Column, ctx: Optional[capi.Context] = None):
    """== "value_counts" (vector_hash.cc): (distinct values in first-occurrence order, int64 counts); the null entry
    counts the nulls (hash_count in both CountOptions modes over the column itself: one of the two is zero per group)."""
    ctx = ctx or default_context()
    from . import columns as cols
    # every row of a group counts, the null group's too (ValueCountsAction::ObserveNullFound, vector_hash.cc:120-163):
    # hash_count over the column read WITHOUT its validity (a binary column through its int32 offsets buffer)
    all_rows = Column(cols.INT32 if values.type == cols.BINARY else values.type, values.data, None, values.length, values.offset, 0)
    out = capi.group_by(ctx, [values], [AggSpec(cols.HASH_COUNT, 0)], [all_rows])
    G = out.num_groups
    return out.keys[0], out.aggs[0][0].data[: G * 8].view(torch.int64)


def DictionaryEncode(values: Column, encode_nulls: bool = False, ctx: Optional[capi.Context] = None):
    """== "dictionary_encode" (vector_hash.cc:165-226, 764-779) for a fixed-width device column: (int32 indices
    Column, dictionary Column).  The dictionary lists the distinct values in first-occurrence order.
    encode_nulls False == DictionaryEncodeOptions::MASK (the default upstream): a null row gets a null index and the
    null has no dictionary entry; True == ENCODE: the null is a dictionary entry like any value."""
    ctx = ctx or default_context()
    from . import columns as cols
    gb = capi.GroupBy(ctx, [values.type], [], [], 0)
    try:
        ids = gb.grouper_consume([values])
        uniq = gb.finalize().keys[0]
    finally:
        gb.close()
    null_group, dictionary = -1, uniq
    if not encode_nulls and uniq.null_count != 0:
        # the one null entry of the dictionary: where is it (G bits to the host), and the dictionary without it
        G = uniq.length
        bits = cols.unpack_bits(uniq.validity[: (G + 7) // 8].cpu().numpy(), G)
        null_group = int(np.flatnonzero(~bits)[0])
        if uniq.type == cols.BINARY:
            # the null entry holds no bytes: the value buffer stays, the int32 offsets lose entry null_group + 1
            offs = uniq.data[: (G + 1) * 4].view(torch.int32)
            kept = torch.cat([offs[: null_group + 1], offs[null_group + 2:]]).contiguous().view(torch.uint8)
            dictionary = Column(uniq.type, kept, None, G - 1, 0, 0, uniq.extra)
        else:
            kept = capi.filter(ctx, uniq, Column(cols.BOOL, uniq.validity, None, G, 0, 0), DROP).col
            dictionary = Column(uniq.type, kept.data, None, kept.length, 0, 0)   # every entry left is valid
    idx, validity = capi.dictionary_indices(ctx, ids, null_group)
    n = values.length
    nulls = 0 if validity is None else (values.null_count if values.null_count >= 0 else -1)
    return Column(cols.INT32, idx.view(torch.uint8), validity, n, 0, nulls), dictionary

