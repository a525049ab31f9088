"""Multi-GPU forms of the hot path: one process per GPU, rows range-sharded.

    dist_filter        no collective: every rank compacts its own row range; the result stays
                       sharded, rank order == row order (a ChunkedArray with one chunk per rank).
    dist_sort_indices  sample sort: local stable radix sort of (key, global row) pairs, splitters
                       from an all-gathered regular sample, one all-to-all of contiguous runs over
                       NVLink, local stable radix sort of the received runs.  Reproduces
                       ChunkedArraySorter's global stable order with NaNs then nulls last
                       (compute/kernels/vector_sort.cc:657-832).
    dist_group_by      low cardinality: every rank aggregates its rows into a partial table (first
                       rows are GLOBAL row numbers), the occupied slots are all-gathered and merged
                       in rank order on every rank (== HashAggregateKernel::merge, which the
                       reference leaves unimplemented, hash_aggregate.cc:795-798).
                       high cardinality: rows are hash-partitioned by key, exchanged with one
                       all-to-all, and each rank aggregates a disjoint key set.

torch.distributed (NCCL on GPUs, gloo in the CPU tests) is the plumbing; all computation goes
through a `LocalOps` object — `CudaOps` (the C ABI) in the product; the tests substitute a numpy
stand-in to exercise the exchange logic with gloo on CPU.
"""
from __future__ import annotations

import os
from typing import List, Optional, Sequence

import torch
import torch.distributed as dist

from .columns import AggSpec, Column

_I64_MIN = -(1 << 63)
_ROW_MASK = (1 << 62) - 1
_TAIL_KEY_BIASED = ((1 << 64) - 2) ^ (1 << 63)   # biased image of the NaN key 2^64-2 (== INT64_MAX - 1)
OVERSAMPLE = 8192   # regular sample per rank: splitter quantile error ~ 1 / sqrt(world * OVERSAMPLE) of the rows


def _bias(keys: torch.Tensor) -> torch.Tensor:
    """uint64 keys held in int64 -> int64 whose signed order is the unsigned key order."""
    return keys ^ _I64_MIN


class CudaOps:
    """LocalOps over the C ABI (arrow1_b200.capi)."""

    def __init__(self, ctx):
        from . import capi
        self.capi, self.ctx = capi, ctx
        self.device = torch.device("cuda:%d" % ctx.device)

    def sort_keys_rows(self, col: Column, order: int, index_base: int):
        return self.capi.sort_keys_rows(self.ctx, col, order, index_base)

    def sort_pairs(self, keys: torch.Tensor, rows: torch.Tensor):
        if keys.numel() == 0:
            return keys, rows
        return self.capi.sort_pairs(self.ctx, keys, rows, want_keys=True)

    def filter(self, values: Column, mask: Column, null_selection: int):
        return self.capi.filter(self.ctx, values, mask, null_selection)

    def groupby(self, key_types, aggs, arg_types, expected_groups):
        return self.capi.GroupBy(self.ctx, key_types, aggs, arg_types, expected_groups)

    def range_partition(self, col: Column, order: int, index_base: int, splitters, n_parts: int):
        return self.capi.sort_range_partition(self.ctx, col, order, index_base, splitters, n_parts)

    # ---- range partition fused with the exchange (peer memory over NVLink)
    def peer_sort_exchange(self, capacity: int, group=None) -> "PeerSortExchange":
        """This rank's receive buffers (>= capacity rows), mapped into every rank; collective (every rank calls
        it with the same capacity)."""
        ex = getattr(self.ctx, "_peer_sort", None)   # lives with the context: CudaOps objects are throw-away
        if ex is None or ex.capacity < capacity:
            if ex is not None:
                ex.close(group)
            ex = PeerSortExchange(self, int(capacity * 1.25) + 4096, group)
            self.ctx._peer_sort = ex
            if not ex.ok:   # no peer mappings on this node (same outcome on every rank): NCCL all-to-all from now on
                self.ctx._peer_sort_failed = True
        return ex

    def range_partition_count(self, col: Column, order: int, splitters, n_parts: int):
        return self.capi.RangePartitionPush(self.ctx, col, order, splitters, n_parts)

    def sort_pairs32(self, keys: torch.Tensor, rows32: torch.Tensor, want_keys: bool):
        if keys.numel() == 0:
            return keys, torch.empty(0, dtype=torch.int64, device=self.device)
        return self.capi.sort_pairs32(self.ctx, keys, rows32, want_keys=want_keys)

    def sort_pairs32_into(self, keys: torch.Tensor, rows32: torch.Tensor, out: torch.Tensor) -> None:
        if keys.numel():
            self.capi.sort_pairs32(self.ctx, keys, rows32, want_keys=False, out=out)

    def sample_keys(self, col: Column, order: int, count: int) -> torch.Tensor:
        """Radix keys of `count` regularly spaced rows (plumbing: a gathered mini-column through b200_sort_keys)."""
        from .columns import BITS
        n = col.length
        if n == 0:
            return torch.empty(0, dtype=torch.int64, device=self.device)
        m = min(count, n)
        pos = ((torch.arange(m, device=self.device, dtype=torch.int64) * 2 + 1) * n) // (2 * m) + col.offset
        width = BITS[col.type] // 8
        vals = col.data[: (col.offset + n) * width].view(-1, width)[pos].contiguous().view(-1)
        validity = None
        nulls = 0
        if col.validity is not None and col.null_count != 0:
            bits = (col.validity[pos >> 3].to(torch.int32) >> (pos & 7).to(torch.int32)) & 1
            nulls = int(m - int(bits.sum().item()))
            padded = torch.zeros(((m + 7) // 8 + 64) * 8, dtype=torch.uint8, device=self.device)
            padded[:m] = bits.to(torch.uint8)
            w = torch.tensor([1, 2, 4, 8, 16, 32, 64, 128], dtype=torch.uint8, device=self.device)
            validity = (padded.view(-1, 8) * w).sum(dim=1, dtype=torch.uint8)
        buf = torch.zeros(vals.numel() + 64, dtype=torch.uint8, device=self.device)
        buf[: vals.numel()] = vals
        mini = Column(col.type, buf, validity, m, 0, nulls)
        keys, _ = self.capi.sort_keys_rows(self.ctx, mini, order, 0)
        return keys

    # ---- hash repartitioning (high-cardinality group-by)
    def partition(self, gb, keys: Sequence[Column], cols: Sequence[Column], n_parts: int, row_base: int):
        """-> (counts[n_parts], [data bytes tensor per col], [validity bytes tensor or None per col], rows)
        in the partitioned layout (destination p contiguous, starting at sum(counts[:p]))."""
        pos, counts = gb.partition_plan(keys, n_parts)
        datas, valids = [], []
        for c in cols:
            d, v = self.capi.partition_scatter(self.ctx, c, pos, c.validity is not None and c.null_count != 0)
            datas.append(d)
            valids.append(v)
        rows = self.capi.partition_rows(self.ctx, keys[0].length, row_base, pos)
        return counts, datas, valids, rows

    def column_from_exchange(self, type_: int, data: torch.Tensor, valid_bytes: Optional[torch.Tensor], n: int) -> Column:
        if valid_bytes is None:
            return Column(type_, data, None, n, 0, 0)
        bitmap, nulls = self.capi.pack_validity(self.ctx, valid_bytes)
        return Column(type_, data, bitmap, n, 0, nulls)


class PeerSortExchange:
    """Receive buffers of the fused range partition + exchange: `capacity` uint64 keys and uint32 rows on
    every rank (b200_peer_alloc), each mapped into every other rank through CUDA IPC so that a rank's scatter
    kernel stores rank q's run directly into rank q's buffers over NVLink."""

    def __init__(self, ops: CudaOps, capacity: int, group=None):
        capi = ops.capi
        self.capacity = capacity
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.keys = self.rows = None
        self.key_ptrs, self.row_ptrs = [], []
        self.ok = False
        self.flag = torch.zeros(1, dtype=torch.int32, device=ops.device)

        def agreed(success: bool) -> bool:
            """Every step that can fail locally (allocation, IPC export / open) is followed by a MIN all-reduce
            of the outcome, so that all ranks take the same path and no collective is left half-entered."""
            t = torch.tensor([1 if success else 0], dtype=torch.int32, device=ops.device)
            dist.all_reduce(t, op=dist.ReduceOp.MIN, group=group)
            return bool(t.item())

        mine = torch.zeros(128, dtype=torch.uint8, device=ops.device)
        try:
            if os.environ.get("B200_NO_PEER"):   # tests: behave as a node without peer mappings
                raise capi.B200Error(9, "peer mappings disabled by B200_NO_PEER")
            self.keys = capi.PeerBuffer(ops.ctx, capacity * 8)
            self.rows = capi.PeerBuffer(ops.ctx, capacity * 4)
            mine = torch.tensor(list(self.keys.handle() + self.rows.handle()), dtype=torch.uint8, device=ops.device)
            good = True
        except capi.B200Error:
            good = False
        if not agreed(good):
            self._release()
            return
        every = torch.empty(self.world * mine.numel(), dtype=torch.uint8, device=ops.device)
        dist.all_gather_into_tensor(every, mine, group=group)
        blob = every.cpu().numpy().tobytes()
        try:
            for r in range(self.world):
                h = blob[r * 128: (r + 1) * 128]
                if r == self.rank:
                    self.key_ptrs.append(self.keys.ptr)
                    self.row_ptrs.append(self.rows.ptr)
                else:
                    self.key_ptrs.append(self.keys.open(h[:64]))
                    self.row_ptrs.append(self.rows.open(h[64:]))
            good = True
        except capi.B200Error:
            good = False
        if not agreed(good):
            self.keys.close_peers()
            self.rows.close_peers()
            dist.barrier(group=group)
            self._release()
            return
        self.ok = True

    def _release(self) -> None:
        for b in (self.keys, self.rows):
            if b is not None:
                b.close()
        self.keys = self.rows = None

    def fence(self, group=None) -> None:
        """Stream-ordered barrier: returns (on the stream) once every rank's earlier work has completed."""
        dist.all_reduce(self.flag, group=group)

    def close(self, group=None) -> None:
        torch.cuda.synchronize()
        dist.barrier(group=group)       # nobody is still writing into a peer
        if self.keys is None:
            return
        self.keys.close_peers()
        self.rows.close_peers()
        dist.barrier(group=group)       # every mapping of this rank's buffers is gone
        self._release()


# ------------------------------------------------------------------------------------------------
def dist_filter(ops, values: Column, mask: Column, null_selection: int = 0):
    """Filter of a row-range shard.  No collective: shard boundaries are arbitrary
    because every bitmap is addressed by (buffer, bit offset)."""
    return ops.filter(values, mask, null_selection)


# ------------------------------------------------------------------------------------------------
def choose_splitters(samples: torch.Tensor, world: int) -> torch.Tensor:
    """`samples`: the biased keys every rank sampled from its sorted run, concatenated (empty
    ranks contribute none).  Returns world-1 splitters; rank j owns biased keys in
    [splitter[j-1], splitter[j]).  Keys >= the NaN key (NaN / null / genuine maxima) all belong to
    the last rank so that one rank can order the three classes (see dist_sort_indices)."""
    if samples.numel() == 0:
        return torch.full((world - 1,), _TAIL_KEY_BIASED, dtype=torch.int64, device=samples.device)
    s, _ = torch.sort(samples)
    pos = (torch.arange(1, world, device=samples.device, dtype=torch.int64) * s.numel()) // world
    spl = s[pos.clamp(max=s.numel() - 1)]
    return torch.clamp(spl, max=_TAIL_KEY_BIASED)


def sample_sorted(biased_sorted: torch.Tensor, count: int) -> torch.Tensor:
    n = biased_sorted.numel()
    if n == 0:
        return biased_sorted[:0]
    m = min(count, n)
    pos = ((torch.arange(m, device=biased_sorted.device, dtype=torch.int64) * 2 + 1) * n) // (2 * m)
    return biased_sorted[pos]


def class_fixup(keys_sorted: torch.Tensor, rows_sorted: torch.Tensor, class_shift: int = 62) -> torch.Tensor:
    """Rows whose key is >= the NaN key: genuine values first (already in key, row order), then
    NaNs, then nulls, each in global row order (vector_sort.cc:269-302).  Returns row numbers
    with the class bits (at `class_shift`) stripped."""
    mask = (1 << class_shift) - 1
    if keys_sorted.numel() == 0:
        return rows_sorted
    # keys are in unsigned order: such rows sit at the very end, so the last key tells whether there are any
    if int(keys_sorted[-1].item()) not in (-1, -2):
        return rows_sorted if class_shift == 30 else rows_sorted & mask   # 32-bit rows: no class bit can be set
    biased = _bias(keys_sorted)
    start = int(torch.searchsorted(biased, torch.tensor([_TAIL_KEY_BIASED], dtype=torch.int64, device=biased.device)).item())
    tail = rows_sorted[start:]
    if tail.numel():
        cls = (tail >> class_shift) & 3
        if bool((cls != 0).any()):
            tail = torch.cat([tail[cls == 0], tail[cls == 1], tail[cls == 2]])
            rows_sorted = torch.cat([rows_sorted[:start], tail])
    return rows_sorted & mask


TRACE: Optional[list] = None   # set to [] to collect (phase name, CUDA event) marks (scripts/dist_sort_breakdown.py)


def _mark(name: str) -> None:
    if TRACE is not None and torch.cuda.is_available():
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        TRACE.append((name, ev))


def dist_sort_indices(ops, col: Column, order: int, row_base: int, group=None):
    """Global SortIndices over the concatenation of every rank's shard (rank order == row order).
    Returns (indices, global_offset): this rank's contiguous slice of the global uint64 index
    array (int64-typed tensor) and where it starts.

    Sample sort with ONE local radix sort per rank: splitters come from a regular sample of every rank's
    (unsorted) keys; each rank range-partitions its (key, global row) pairs by the splitters with a STABLE
    partition kernel, the runs reach their destination rank, and since the runs a rank receives are in
    global row order when concatenated in source-rank order, one stable sort by key yields the (key, row)
    order.  Keys >= the NaN key all go to the last rank, which orders values / NaNs / nulls (class_fixup).

    On GPUs the partition's scatter IS the exchange (`_exchange_fused`): it stores every run straight into
    the destination rank's receive buffer through NVLink peer mappings.  The NCCL all-to-all form
    (`_exchange_all_to_all`) remains for row numbers beyond 2^30 and for the gloo tests."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    if world == 1:
        keys, rows = ops.sort_keys_rows(col, order, row_base)
        ks, rs = ops.sort_pairs(keys, rows)
        return class_fixup(ks, rs), 0
    dev = ops.device
    _mark("start")
    # ---- splitters from a regular sample of every rank's keys
    got = _bias(ops.sample_keys(col, order, OVERSAMPLE))
    W = OVERSAMPLE + 3                      # the sample, then (sample count, rows, row base): one all-gather
    smp = torch.zeros(W, dtype=torch.int64, device=dev)
    smp[: got.numel()] = got
    smp[OVERSAMPLE:] = torch.tensor([got.numel(), col.length, row_base], dtype=torch.int64)
    all_smp = torch.empty(world * W, dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(all_smp, smp, group=group)
    all_smp = all_smp.view(world, W)
    metas = all_smp[:, OVERSAMPLE:].tolist()
    valid = torch.cat([all_smp[r, : metas[r][0]] for r in range(world)])
    splitters = choose_splitters(valid, world)
    spl_unsigned = [(int(x) ^ _I64_MIN) & ((1 << 64) - 1) for x in splitters.tolist()]
    _mark("splitters")
    fused = (hasattr(ops, "peer_sort_exchange") and FUSED_EXCHANGE and world < 16
             and not getattr(getattr(ops, "ctx", None), "_peer_sort_failed", False)
             and max(m[2] + m[1] for m in metas) <= (1 << 30))
    if fused:
        return _exchange_fused(ops, col, order, row_base, spl_unsigned, max(m[1] for m in metas), world, rank, group)
    return _exchange_all_to_all(ops, col, order, row_base, spl_unsigned, world, rank, group)


FUSED_EXCHANGE = True   # tests flip this to run the NCCL all-to-all form on GPUs too


_NAN_KEY = (1 << 64) - 2


def _exchange_fused(ops, col, order, row_base, spl_unsigned, max_rows, world, rank, group):
    """count -> all-gather of the count matrix -> scatter into the peers' receive buffers -> fence -> one sort.
    The rows whose key is >= the NaN key (NaN, null, genuine maxima) form one extra partition that also
    goes to the last rank, behind its other runs, so that their number is known from the counts alone."""
    dev = ops.device
    parts = world + 1
    plan = ops.range_partition_count(col, order, list(spl_unsigned) + [_NAN_KEY], parts)
    mine = torch.tensor(plan.counts, dtype=torch.int64, device=dev)
    matrix = torch.empty(world * parts, dtype=torch.int64, device=dev)
    # also orders the exchange: it completes here only after every rank has finished the sort that read its
    # receive buffer in the previous call (their all-gather is enqueued behind it)
    dist.all_gather_into_tensor(matrix, mine, group=group)
    m = matrix.view(world, parts).tolist()                     # m[src][partition]
    body = [sum(m[src][q] for src in range(world)) for q in range(world)]
    tail = sum(m[src][world] for src in range(world))
    incoming = body[:-1] + [body[-1] + tail]
    _mark("count")
    ex = ops.peer_sort_exchange(max(max(incoming), max_rows), group)
    if not ex.ok:
        return _exchange_all_to_all(ops, col, order, row_base, spl_unsigned, world, rank, group)
    offsets = [sum(m[src][q] for src in range(rank)) for q in range(world)]
    offsets.append(body[-1] + sum(m[src][world] for src in range(rank)))
    plan.push(row_base, ex.key_ptrs + [ex.key_ptrs[-1]], ex.row_ptrs + [ex.row_ptrs[-1]], offsets, rows32=True)
    ex.fence(group)                                            # every run addressed to this rank has landed
    _mark("push")
    total = incoming[rank]
    rk = ex.keys.tensor(torch.int64, total)
    rr = ex.rows.tensor(torch.int32, total)
    if rank == world - 1 and tail and hasattr(ops, "sort_pairs32_into"):
        # the tail partition (keys >= the NaN key) sits behind the other runs and its size is known from the counts:
        # it is sorted on its own — thousands of equal keys — so that the rest, keys that nearly identify their
        # rows, can take the sort's upper-key-word-first plan (sort.cu HwState)
        out = torch.empty(max(total, 1), dtype=torch.int64, device=dev)[:total]
        body = total - tail
        if body:
            ops.sort_pairs32_into(rk[:body], rr[:body], out[:body])
        ops.sort_pairs32_into(rk[body:], rr[body:], out[body:])
    else:
        _, out = ops.sort_pairs32(rk, rr, want_keys=False)
    _mark("sort_pairs")
    if rank == world - 1 and tail:
        # rows arrive as global row | class << 30; the stable sort left the tail partition at the end in
        # (key, row) order: genuine maxima, then NaNs, then nulls (vector_sort.cc:269-302)
        t = out[total - tail:]
        cls = (t >> 30) & 3
        if bool((cls != 0).any()):
            out[total - tail:] = torch.cat([t[cls == 0], t[cls == 1], t[cls == 2]]) & ((1 << 30) - 1)
    _mark("class_fixup")
    return out, sum(incoming[:rank])


def _exchange_all_to_all(ops, col, order, row_base, spl_unsigned, world, rank, group):
    dev = ops.device
    # ---- stable range partition, then one all-to-all of keys and one of rows
    ks, rs, send_counts = ops.range_partition(col, order, row_base, spl_unsigned, world)
    _mark("range_partition")
    sc = torch.tensor(send_counts, dtype=torch.int64, device=dev)
    rc = torch.empty(world, dtype=torch.int64, device=dev)
    dist.all_to_all_single(rc, sc, group=group)
    recv_counts = rc.tolist()
    total = sum(recv_counts)
    rk = torch.empty(total, dtype=torch.int64, device=dev)
    rr = torch.empty(total, dtype=torch.int64, device=dev)
    dist.all_to_all_single(rk, ks.contiguous(), recv_counts, send_counts, group=group)
    dist.all_to_all_single(rr, rs.contiguous(), recv_counts, send_counts, group=group)
    _mark("all_to_all")
    # ---- received runs, concatenated in source-rank order, are in global row order: a stable sort by key
    ks2, rs2 = ops.sort_pairs(rk, rr)
    _mark("sort_pairs")
    out = class_fixup(ks2, rs2) if rank == world - 1 else (rs2 & _ROW_MASK)
    _mark("class_fixup")
    tot = torch.tensor([total], dtype=torch.int64, device=dev)
    all_tot = torch.empty(world, dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(all_tot, tot, group=group)
    offset = int(all_tot[:rank].sum().item())
    _mark("offsets")
    return out, offset


# ------------------------------------------------------------------------------------------------
def dist_group_by(ops, keys: Sequence[Column], aggs: Sequence[AggSpec], args: Sequence[Column], row_base: int,
                  expected_groups: int = 0, group=None):
    """Low-cardinality grouped aggregation: local partial table -> all-gather of the occupied
    slots -> merge in rank order on every rank (bitwise identical results on all ranks; float
    sums within the stated tolerance of the single-pass result).  Returns the finalized output of
    the merged table (every rank holds the full result, in the reference's first-occurrence
    order because first rows are global)."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    key_types = [k.type for k in keys]
    arg_types = [a.type for a in args]
    part = ops.groupby(key_types, aggs, arg_types, expected_groups)
    try:
        part.consume(keys, args, row_base)
        if world == 1:
            return part.finalize()
        state = part.export_state()
        sb = part.slot_bytes()
        dev = state.device
        n = torch.tensor([state.numel()], dtype=torch.int64, device=dev)
        all_n = torch.empty(world, dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(all_n, n, group=group)
        sizes = all_n.tolist()
        cap = max(max(sizes), sb)
        padded = torch.zeros(cap, dtype=torch.uint8, device=dev)
        padded[: state.numel()] = state
        gathered = torch.empty(world * cap, dtype=torch.uint8, device=dev)
        dist.all_gather_into_tensor(gathered, padded, group=group)
        merged = ops.groupby(key_types, aggs, arg_types, max(expected_groups, max(sizes) // sb))
        try:
            for r in range(world):
                if sizes[r]:
                    merged.import_state(gathered[r * cap: r * cap + sizes[r]])
            return merged.finalize()
        finally:
            merged.close()
    finally:
        part.close()


# ------------------------------------------------------------------------------------------------
class ShardedGroups:
    """Result of the repartitioned group-by on one rank: this rank's DISJOINT set of groups (finalized
    output in local first-occurrence order) plus each group's global first row, which ranks the groups
    of all ranks into the reference's order (hash_aggregate.cc:391-403)."""

    def __init__(self, out, first_rows, stats=None):
        self.out, self.first_rows = out, first_rows
        self.stats = stats or {}   # which way the local aggregation went (b200_groupby_stat)

    @property
    def num_groups(self):
        return self.out.num_groups


def _all_to_all_bytes(t: torch.Tensor, width: int, send_counts, recv_counts, group):
    """all-to-all of a byte tensor holding `width`-byte elements, counts in elements."""
    total = sum(recv_counts)
    out = torch.empty(max(total, 1) * width, dtype=torch.uint8, device=t.device)[: total * width]
    dist.all_to_all_single(out, t.contiguous(), [c * width for c in recv_counts], [c * width for c in send_counts], group=group)
    return out


def dist_group_by_repartition(ops, keys: Sequence[Column], aggs: Sequence[AggSpec], args: Sequence[Column], row_base: int,
                              expected_groups: int = 0, group=None) -> ShardedGroups:
    """High-cardinality grouped aggregation: rows are hash-partitioned by key
    on every rank, exchanged with one all-to-all per column (+ the global row numbers) over NVLink, and
    each rank aggregates the DISJOINT key set it owns.  Every key's rows meet on one rank, so counts,
    min/max and integer sums are exact and a float sum sees all of its addends on one GPU.  Results stay
    sharded; `first_rows` restores the reference's global group order when gathered."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    key_types = [k.type for k in keys]
    arg_types = [a.type for a in args]
    n = keys[0].length
    # distinct columns are exchanged once
    cols, index = [], {}
    for c in list(keys) + list(args):
        ident = (id(c.data), id(c.validity), c.offset)
        if ident not in index:
            index[ident] = len(cols)
            cols.append(c)
    gb = ops.groupby(key_types, aggs, arg_types, max(expected_groups // world, 1) if expected_groups else 0)
    try:
        if world == 1:
            gb.consume(keys, args, row_base)
            return ShardedGroups(gb.finalize(), gb.first_rows())
        counts, datas, valids, rows = ops.partition(gb, keys, cols, world, row_base)
        dev = rows.device
        sc = torch.tensor(counts, dtype=torch.int64, device=dev)
        rc = torch.empty(world, dtype=torch.int64, device=dev)
        dist.all_to_all_single(rc, sc, group=group)
        recv_counts = rc.tolist()
        total = sum(recv_counts)
        # whether a column carries validity must agree on every rank
        has_valid = torch.tensor([1 if v is not None else 0 for v in valids], dtype=torch.int64, device=dev)
        dist.all_reduce(has_valid, op=dist.ReduceOp.MAX, group=group)
        has_valid = has_valid.tolist()
        recv_cols = []
        for c, d, v, hv in zip(cols, datas, valids, has_valid):
            width = _WIDTH_OF(c)
            rd = _all_to_all_bytes(d, width, counts, recv_counts, group)
            rv = None
            if hv:
                if v is None:
                    v = torch.ones(n, dtype=torch.uint8, device=dev)
                rv = _all_to_all_bytes(v, 1, counts, recv_counts, group)
            recv_cols.append(ops.column_from_exchange(c.type, rd, rv, total))
        rrows = _all_to_all_bytes(rows.view(torch.uint8), 8, counts, recv_counts, group).view(torch.int64)
        rkeys = [recv_cols[index[(id(c.data), id(c.validity), c.offset)]] for c in keys]
        rargs = [recv_cols[index[(id(c.data), id(c.validity), c.offset)]] for c in args]
        # the send-side copies are dead; the partition-then-aggregate path of the local aggregation wants 48 B per
        # received row of scratch from cudaMalloc, which PyTorch's cached blocks may be sitting on
        del datas, valids, rows, d, v
        if dev.type == "cuda" and torch.cuda.mem_get_info(dev)[0] < total * 48 + (2 << 30):
            torch.cuda.empty_cache()
        if total:
            gb.consume_rows(rkeys, rargs, rrows)
        stats = {}
        if hasattr(gb, "stat"):
            stats = {k: gb.stat(k) for k in ("partitioned_calls", "partitioned_alloc_failures", "partitioned_span_refusals")}
        return ShardedGroups(gb.finalize(), gb.first_rows(), stats)
    finally:
        gb.close()


def _WIDTH_OF(c: Column) -> int:
    from .columns import BITS
    return max(BITS[c.type] // 8, 1)


# ------------------------------------------------------------------------------------------------
# bench.py hooks: one step of a workload over this rank's shard

def _rank_world():
    return (dist.get_rank(), dist.get_world_size()) if dist.is_initialized() else (0, 1)


def bench_step(ctx, wl, cols_, bufs):
    ops = CudaOps(ctx)
    rank, _ = _rank_world()
    if wl.name.startswith("sort"):
        return dist_sort_indices(ops, cols_[0], 0, rank * wl.n)
    if wl.name.startswith("filter"):
        return wl.step(ctx, cols_, bufs)
    if wl.name == "groupby_c3":
        k, v = cols_
        return dist_group_by(ops, [k], wl._specs(), [v] * len(wl.agg_names), rank * wl.n, wl.n_keys)
    if wl.name == "groupby_c4":
        k, v = cols_
        return dist_group_by_repartition(ops, [k], wl._specs(), [v] * len(wl.agg_names), rank * wl.n, wl.n_keys)
    raise SystemExit("bench: workload %s has no multi-GPU form yet" % wl.name)


def bench_self_check(ctx, wl, cols_, bufs):
    """One more step of the workload, outside the timed region, checked against what the counter-based generator
    says (every SCALE run doubles as a parity run).  Collective: every rank calls it.  Returns a JSON-able dict.
      sort      this rank's slice of the global index array is strictly increasing in (class, key, row) with the keys
                REGENERATED from the row numbers (rows come from every rank), the boundary to the next rank's slice
                is in order too, all rows lie in [0, N), and the slices hold N rows together  ==> a permutation in
                the reference's stable order
      filter    output length == number of set mask bits of the shard
      group-by  low cardinality: every rank holds the same number of groups and the counts add up to the valid rows
                of all shards; high cardinality: the per-rank counts (sum / mean) add up to all rows"""
    import numpy as np
    from . import synth
    rank, world = _rank_world()
    res = bench_step(ctx, wl, cols_, bufs)
    dev = cols_[0].data.device
    n_total = wl.n * world

    def allsum(x):
        t = torch.tensor([int(x)], dtype=torch.int64, device=dev)
        if world > 1:
            dist.all_reduce(t)
        return int(t.item())

    if wl.name.startswith("sort"):
        out, off = res
        idx = out.view(torch.int64)
        m = idx.numel()
        ok = allsum(m) == n_total and bool(((idx >= 0) & (idx < n_total)).all()) if m else allsum(0) == n_total
        valid = (synth._lsr(synth.r_at(6, idx), 33) % 100) != 0
        if wl.name == "sort_i64":
            key = synth.r_at(4, idx)
        else:
            key = (synth._lsr(synth.r_at(5, idx), 11).to(torch.float64) * 2.0 ** -53 * 1e14).view(torch.int64)   # positive doubles order like their bits
        cls = (~valid).to(torch.int64)
        key = torch.where(valid, key, torch.zeros_like(key))
        if m > 1:
            a = (cls[:-1], key[:-1], idx[:-1])
            b = (cls[1:], key[1:], idx[1:])
            lt = (a[0] < b[0]) | ((a[0] == b[0]) & ((a[1] < b[1]) | ((a[1] == b[1]) & (a[2] < b[2]))))
            ok = ok and bool(lt.all())
        # boundaries: (first, last) triple of every rank
        edge = torch.zeros(6, dtype=torch.int64, device=dev)
        if m:
            edge = torch.stack([cls[0], key[0], idx[0], cls[-1], key[-1], idx[-1]])
        edges = [torch.zeros_like(edge) for _ in range(world)]
        sizes = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(world)]
        if world > 1:
            dist.all_gather(edges, edge)
            dist.all_gather(sizes, torch.tensor([m], dtype=torch.int64, device=dev))
        else:
            edges, sizes = [edge], [torch.tensor([m], device=dev)]
        prev = None
        for e, sz in zip(edges, sizes):
            if int(sz.item()) == 0:
                continue
            e = e.tolist()
            if prev is not None and not (tuple(prev) < tuple(e[:3])):
                ok = False
            prev = e[3:]
        return {"workload": wl.name, "rows_total": n_total, "permutation_in_stable_order": bool(allsum(0 if ok else 1) == 0)}
    if wl.name.startswith("filter"):
        from .capi import FilterPlan
        mask = cols_[1]
        bits = mask.data[: (mask.length + 7) // 8]
        w = torch.tensor([1, 2, 4, 8, 16, 32, 64, 128], dtype=torch.uint8, device=dev)
        want = int(((bits.unsqueeze(1) & w) != 0).reshape(-1)[: mask.length].sum().item())
        ctx.synchronize()
        got = res.length if res.length >= 0 else int(res.scratch.length_tensor().item())   # the bench step keeps it on the device
        return {"workload": wl.name, "rows_total": n_total, "selected_rows_total": allsum(got),
                "output_length_matches_mask": bool(allsum(0 if got == want else 1) == 0)}
    if wl.name == "groupby_c3":
        counts = res.aggs[1][0].values()
        valid_local = wl.n - cols_[1].null_count
        total_valid = allsum(valid_local)
        same = allsum(res.num_groups) == res.num_groups * world
        return {"workload": wl.name, "rows_total": n_total, "groups": int(res.num_groups),
                "counts_add_up_to_valid_rows": bool(int(counts.sum()) == total_valid and same)}
    if wl.name == "groupby_c4":
        out = res.out
        G = out.num_groups
        sums = out.aggs[0][0].data[: G * 8].view(torch.float64)
        means = out.aggs[1][0].data[: G * 8].view(torch.float64)
        local = int(torch.round(sums / means).sum().item()) if G else 0
        return {"workload": wl.name, "rows_total": n_total, "groups_total": allsum(G),
                "counts_add_up_to_rows": bool(allsum(local) == n_total), "local_aggregation_rank0": dict(res.stats)}
    return {"workload": wl.name, "checked": False}


def SortIndicesSharded(values: Column, order: int = 0, row_base: int = 0, ctx=None, _wait: bool = True, group=None):
    """The multi-GPU SortIndices from this rank's HOST shard: staged to the device, sorted across the ranks
    (dist_sort_indices), this rank's slice of the global index array read back.  Returns (uint64 numpy indices,
    global offset of the slice).  Collective: every rank calls it, in the same order.  Takes the `ctx` / `_wait`
    keywords of compute.Pipeline lanes, so several calls can be in flight (their collectives are issued in
    the same order on every rank; each lane's context owns its own exchange buffers)."""
    from . import compute
    import numpy as np
    ctx = ctx or compute.default_context()
    dv, b0 = compute._h2d(values, ctx)   # already on the device when the caller staged it (compute.Pipeline.stage)
    out, offset = dist_sort_indices(CudaOps(ctx), dv, order, row_base, group)
    h = compute._d2h(out, out.numel() * 8)
    if _wait:
        ctx.stream.synchronize()
    compute.Transfer.h2d, compute.Transfer.d2h = b0, out.numel() * 8
    return h.view(np.uint64), offset


SortIndicesSharded._b200_async = True


def bench_e2e_step(ctx, wl, host_cols):
    """Host shard -> device -> multi-GPU op -> host (the result slice / merged groups)."""
    from . import compute
    ops = CudaOps(ctx)
    rank, _ = _rank_world()
    if wl.name.startswith("sort"):
        return SortIndicesSharded(host_cols[0], 0, rank * wl.n, ctx=ctx)
    dev_cols, nbytes = [], 0
    for c in host_cols:
        d, b = compute._h2d(c, ctx)
        dev_cols.append(d)
        nbytes += b
    compute.Transfer.h2d = nbytes
    if wl.name.startswith("filter"):
        return wl.e2e_step(host_cols)
    if wl.name == "groupby_c3":
        k, v = dev_cols
        res = dist_group_by(ops, [k], wl._specs(), [v] * len(wl.agg_names), rank * wl.n, wl.n_keys)
        d2h = 0
        for cols__ in res.aggs:
            for c in cols__:
                d2h += c.to_host().nbytes()
        compute.Transfer.d2h = d2h
        return res
    if wl.name == "groupby_c4":
        k, v = dev_cols
        res = dist_group_by_repartition(ops, [k], wl._specs(), [v] * len(wl.agg_names), rank * wl.n, wl.n_keys)
        d2h = 0
        for cols__ in res.out.aggs + [res.out.keys]:
            for c in cols__:
                d2h += c.to_host().nbytes()
        compute.Transfer.d2h = d2h
        return res
    raise SystemExit("bench: workload %s has no multi-GPU form yet" % wl.name)
