"""World-size-2/3 gloo runs of the multi-GPU exchange logic (arrow1_b200/dist.py) on CPU.

The CUDA kernels cannot run here, so the per-rank computation is a numpy stand-in (`NumpyOps`)
with the C ABI's contracts (b200_sort_keys_rows / b200_sort_pairs / groupby export-import);
what is under test is the host logic: splitter choice, run exchange, class fix-up, slice offsets,
partial-table merge order.  The global result is compared with the oracle on the whole column."""
import os
import pickle
import socket
import sys
import tempfile

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import oracle  # noqa: E402
from arrow1_b200 import columns as cols  # noqa: E402
from arrow1_b200.columns import AggSpec, Column  # noqa: E402

U64 = np.uint64
NAN_KEY, NULL_KEY = U64(2**64 - 2), U64(2**64 - 1)


def radix_keys(col: Column, order: int):
    """numpy image of sort.cu's transform_key + class (0 value, 1 NaN, 2 null)."""
    v = col.values()
    valid = col.valid()
    dt = v.dtype
    bits = dt.itemsize * 8
    wmask = U64(2**bits - 1)
    sign = U64(1 << (bits - 1))
    raw = v.view({1: np.uint8, 2: np.uint16, 4: np.uint32, 8: np.uint64}[dt.itemsize]).astype(U64)
    cls = np.zeros(len(v), dtype=np.uint8)
    if dt.kind == "i":
        k = raw ^ sign
    elif dt.kind == "f":
        isnan = np.isnan(v)
        cls[isnan] = 1
        k = raw.copy()
        k[k == sign] = U64(0)
        neg = (k & sign) != 0
        k = np.where(neg, ~k & wmask, k | sign)
    else:
        k = raw
    if order:
        k = k ^ wmask
    cls[~valid] = 2
    k = np.where(cls == 1, NAN_KEY, k)
    k = np.where(cls == 2, NULL_KEY, k)
    return k.astype(U64), cls


class NumpyGroupBy:
    """Partial aggregation table with opaque export/import (sum, count, min, max per key)."""
    SLOT = 48

    def __init__(self):
        self.t = {}

    def consume(self, keys, args, row_base):
        k = keys[0].values()
        kv = keys[0].valid()
        v = args[0].values().astype(np.float64)
        vv = args[0].valid()
        for i in range(len(k)):
            key = (int(k[i]), bool(kv[i]))
            s = self.t.setdefault(key, [row_base + i, 0.0, 0, np.inf, -np.inf])
            s[0] = min(s[0], row_base + i)
            if vv[i]:
                s[1] += v[i]
                s[2] += 1
                s[3] = min(s[3], v[i])
                s[4] = max(s[4], v[i])

    def consume_rows(self, keys, args, row_ids):
        rows = row_ids.numpy()
        k = keys[0].values()
        kv = keys[0].valid()
        v = args[0].values().astype(np.float64)
        vv = args[0].valid()
        for i in range(len(k)):
            key = (int(k[i]) if kv[i] else 0, bool(kv[i]))
            s = self.t.setdefault(key, [int(rows[i]), 0.0, 0, np.inf, -np.inf])
            s[0] = min(s[0], int(rows[i]))
            if vv[i]:
                s[1] += v[i]
                s[2] += 1
                s[3] = min(s[3], v[i])
                s[4] = max(s[4], v[i])

    def first_rows(self):
        return torch.tensor(sorted(s[0] for s in self.t.values()), dtype=torch.int64)

    def partition_plan(self, keys, n_parts):
        """numpy image of b200_groupby_partition_plan: any deterministic key hash will do here (what is
        under test is the exchange logic); rows of one destination contiguous."""
        k = keys[0].values().astype(np.int64)
        kv = keys[0].valid()
        dest = np.where(kv, (k * 2654435761 >> 7) % n_parts, 0).astype(np.int64)
        order = np.argsort(dest, kind="stable")
        pos = np.empty(len(k), dtype=np.int64)
        pos[order] = np.arange(len(k))
        return torch.from_numpy(pos), [int((dest == p).sum()) for p in range(n_parts)]

    def slot_bytes(self):
        return self.SLOT

    def export_state(self):
        out = np.zeros((len(self.t), 6), dtype=np.float64)
        for j, ((key, kvalid), s) in enumerate(self.t.items()):
            out[j] = [key if kvalid else -1, s[0], s[1], s[2], s[3], s[4]]
        return torch.from_numpy(out.view(np.uint8).reshape(-1).copy())

    def import_state(self, state):
        a = state.numpy().view(np.float64).reshape(-1, 6)
        for row in a:
            key = (int(row[0]), True) if row[0] >= 0 else (0, False)
            s = self.t.setdefault(key, [int(row[1]), 0.0, 0, np.inf, -np.inf])
            s[0] = min(s[0], int(row[1]))
            s[1] += row[2]
            s[2] += int(row[3])
            s[3] = min(s[3], row[4])
            s[4] = max(s[4], row[5])

    def finalize(self):
        items = sorted(self.t.items(), key=lambda kv: kv[1][0])
        return [(k, s[1], s[2], s[3], s[4]) for k, s in items]

    def close(self):
        pass


class NumpyOps:
    device = torch.device("cpu")

    def sort_keys_rows(self, col, order, index_base):
        k, cls = radix_keys(col, order)
        rows = (np.arange(len(k), dtype=U64) + U64(index_base)) | (cls.astype(U64) << U64(62))
        return torch.from_numpy(k.view(np.int64).copy()), torch.from_numpy(rows.view(np.int64).copy())

    def sort_pairs(self, keys, rows):
        k = keys.numpy().view(U64)
        p = np.argsort(k, kind="stable")
        return torch.from_numpy(k[p].view(np.int64).copy()), torch.from_numpy(rows.numpy()[p].copy())

    def filter(self, values, mask, null_selection):
        return oracle.filter(values, mask, null_selection, kind="port")

    def sample_keys(self, col, order, count):
        k, _ = radix_keys(col, order)
        n = len(k)
        if n == 0:
            return torch.empty(0, dtype=torch.int64)
        m = min(count, n)
        pos = ((np.arange(m, dtype=np.int64) * 2 + 1) * n) // (2 * m)
        return torch.from_numpy(k[pos].view(np.int64).copy())

    def range_partition(self, col, order, index_base, splitters, n_parts):
        """numpy image of b200_sort_range_partition: dest = number of splitters <= key, stable."""
        k, cls = radix_keys(col, order)
        rows = (np.arange(len(k), dtype=U64) + U64(index_base)) | (cls.astype(U64) << U64(62))
        spl = np.array(splitters, dtype=U64)
        dest = np.searchsorted(spl, k, side="right") if len(spl) else np.zeros(len(k), dtype=np.int64)
        p = np.argsort(dest, kind="stable")
        return (torch.from_numpy(k[p].view(np.int64).copy()), torch.from_numpy(rows[p].view(np.int64).copy()),
                [int((dest == q).sum()) for q in range(n_parts)])

    def groupby(self, key_types, aggs, arg_types, expected_groups):
        return NumpyGroupBy()

    def partition(self, gb, keys, cols_, n_parts, row_base):
        pos, counts = gb.partition_plan(keys, n_parts)
        p = pos.numpy()
        datas, valids = [], []
        for c in cols_:
            v = c.values()
            out = np.empty_like(v)
            out[p] = v
            datas.append(torch.from_numpy(out.view(np.uint8).reshape(-1).copy()))
            if c.validity is not None and c.null_count != 0:
                vb = np.empty(len(v), dtype=np.uint8)
                vb[p] = c.valid().astype(np.uint8)
                valids.append(torch.from_numpy(vb))
            else:
                valids.append(None)
        rows = np.empty(len(p), dtype=np.int64)
        rows[p] = np.arange(len(p), dtype=np.int64) + row_base
        return counts, datas, valids, torch.from_numpy(rows)

    def column_from_exchange(self, type_, data, valid_bytes, n):
        from arrow1_b200.columns import NP_OF
        v = data.numpy().view(NP_OF[type_])[:n].copy()
        return Column.from_numpy(v, None if valid_bytes is None else valid_bytes.numpy().astype(bool), type=type_)


def shard(col: Column, world: int):
    """Row-range shards (uneven on purpose); returns [(column, row_base)]."""
    n = col.length
    v, valid = col.values(), col.valid()
    bounds = [0] + [n * (r + 1) // world + (3 if r % 2 == 0 and n > 10 else 0) for r in range(world - 1)] + [n]
    bounds = [min(b, n) for b in bounds]
    out = []
    for r in range(world):
        lo, hi = bounds[r], bounds[r + 1]
        c = Column.from_numpy(v[lo:hi].copy(), valid[lo:hi].copy() if col.validity is not None else None, type=col.type)
        out.append((c, lo))
    return out


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, job, path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from arrow1_b200 import dist as bd
    ops = NumpyOps()
    res = []
    for case in job:
        kind = case["kind"]
        if kind == "sort":
            col, base = case["shards"][rank]
            out, off = bd.dist_sort_indices(ops, col, case["order"], base)
            res.append((out.numpy().view(U64).copy(), off))
        elif kind == "groupby":
            (k, v), base = case["shards"][rank]
            res.append(bd.dist_group_by(ops, [k], [AggSpec(cols.HASH_SUM)], [v], base))
        elif kind == "groupby_repart":
            (k, v), base = case["shards"][rank]
            sg = bd.dist_group_by_repartition(ops, [k], [AggSpec(cols.HASH_SUM)], [v], base)
            res.append((sg.out, sg.first_rows.numpy().copy()))
        elif kind == "filter":
            (v, m), base = case["shards"][rank]
            r = bd.dist_filter(ops, v, m, case["null_selection"])
            res.append((r.values().copy(), r.valid().copy(), r.length))
    with open(path + ".%d" % rank, "wb") as f:
        pickle.dump(res, f)
    dist.barrier()
    dist.destroy_process_group()


def run_world(world, job):
    port = _free_port()
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "out")
        mp.spawn(_worker, args=(world, port, job, path), nprocs=world, join=True)
        return [pickle.load(open(path + ".%d" % r, "rb")) for r in range(world)]


def _sort_cases(rng):
    cases = []
    n = 5000
    i64 = rng.integers(-2**63, 2**63 - 1, size=n, dtype=np.int64)
    i64[rng.integers(0, n, 40)] = np.iinfo(np.int64).max     # genuine maxima next to nulls
    i64[rng.integers(0, n, 40)] = np.iinfo(np.int64).min
    cases.append(Column.from_numpy(i64, rng.random(n) > 0.05))
    f = rng.random(n) * 2e3 - 1e3
    for val in (np.nan, -0.0, 0.0, np.inf, -np.inf):
        f[rng.integers(0, n, 60)] = val
    cases.append(Column.from_numpy(f, rng.random(n) > 0.1))
    dup = rng.integers(0, 7, size=n).astype(np.int32)          # heavy duplicates: stability across ranks
    cases.append(Column.from_numpy(dup, rng.random(n) > 0.3))
    cases.append(Column.from_numpy(np.zeros(n, dtype=np.int64), rng.random(n) > 2.0))   # all null
    cases.append(Column.from_numpy(rng.integers(0, 255, size=17, dtype=np.uint8)))      # tiny: empty ranks' samples
    u64 = rng.integers(0, 2**64 - 1, size=n, dtype=np.uint64)
    u64[rng.integers(0, n, 50)] = U64(2**64 - 1)
    u64[rng.integers(0, n, 50)] = U64(2**64 - 2)
    cases.append(Column.from_numpy(u64, rng.random(n) > 0.05))
    return cases


@pytest.mark.parametrize("world", [2, 3])
def test_dist_sort_indices_matches_oracle(world):
    rng = np.random.default_rng(100 + world)
    job, wants = [], []
    for col in _sort_cases(rng):
        for order in (0, 1):
            job.append({"kind": "sort", "order": order, "shards": shard(col, world)})
            wants.append(oracle.sort_indices(col, order, kind="port"))
    results = run_world(world, job)
    for ci, want in enumerate(wants):
        parts = [results[r][ci] for r in range(world)]
        offs = [p[1] for p in parts]
        assert offs == list(np.cumsum([0] + [len(p[0]) for p in parts[:-1]])), "slice offsets"
        got = np.concatenate([p[0] for p in parts])
        np.testing.assert_array_equal(got, want, err_msg="case %d" % ci)


def test_dist_group_by_and_filter_world2():
    world = 2
    rng = np.random.default_rng(5)
    n = 3000
    keys = Column.from_numpy(rng.integers(0, 50, size=n).astype(np.int32), rng.random(n) > 0.02)
    vals = Column.from_numpy(rng.integers(0, 1000, size=n).astype(np.float64), rng.random(n) > 0.1)
    ks, vs = shard(keys, world), shard(vals, world)
    gshards = [((ks[r][0], vs[r][0]), ks[r][1]) for r in range(world)]
    mask = Column.from_numpy(rng.random(n) < 0.5, rng.random(n) > 0.1)
    ms = shard(mask, world)
    fshards = [((vs[r][0], ms[r][0]), vs[r][1]) for r in range(world)]
    job = [{"kind": "groupby", "shards": gshards}, {"kind": "filter", "shards": fshards, "null_selection": 1}]
    results = run_world(world, job)
    want = oracle.groupby([keys], [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_COUNT)], [vals, vals], kind="port")
    for r in range(world):
        got = results[r][0]
        assert len(got) == want.num_groups
        wk, wkv = want.keys[0]
        for g, (key, s, c, mn, mx) in enumerate(got):
            assert key[1] == bool(wkv[g]), "null-key group position"
            if key[1]:
                assert key[0] == int(wk[g]), "first-occurrence group order must be global"
            assert c == int(want.aggs[1][0][0][g])
            if c:
                assert s == want.aggs[0][0][0][g]   # integer-valued doubles: exact in any order
    # filter: concatenation of the shard outputs == the whole-array result
    wf = oracle.filter(vals, mask, 1, kind="port")
    gv = np.concatenate([results[r][1][0] for r in range(world)])
    gvalid = np.concatenate([results[r][1][1] for r in range(world)])
    assert sum(results[r][1][2] for r in range(world)) == wf.length
    np.testing.assert_array_equal(gvalid, wf.valid())
    np.testing.assert_array_equal(gv[gvalid], wf.values()[wf.valid()])


@pytest.mark.parametrize("world", [2, 3])
def test_dist_group_by_repartition_matches_oracle(world):
    """High-cardinality form: hash partition -> all-to-all of (key, value, validity, global row) ->
    disjoint local aggregation.  Gathering every rank's groups and ranking them by global first row
    must give the oracle's output exactly (group order, null-key group, counts, sums)."""
    rng = np.random.default_rng(40 + world)
    n = 4000
    keys = Column.from_numpy(rng.integers(-2000, 2000, size=n).astype(np.int64), rng.random(n) > 0.01)
    vals = Column.from_numpy(rng.integers(0, 1000, size=n).astype(np.float64), rng.random(n) > 0.1)
    ks, vs = shard(keys, world), shard(vals, world)
    gshards = [((ks[r][0], vs[r][0]), ks[r][1]) for r in range(world)]
    results = run_world(world, [{"kind": "groupby_repart", "shards": gshards}])
    want = oracle.groupby([keys], [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_COUNT)], [vals, vals], kind="port")
    merged, seen = [], set()
    for r in range(world):
        groups, first_rows = results[r][0]
        assert len(groups) == len(first_rows)
        for (key, s, c, mn, mx), fr in zip(groups, first_rows):
            assert key not in seen, "key sets of the ranks must be disjoint"
            seen.add(key)
            merged.append((int(fr), key, s, c))
    merged.sort()
    assert len(merged) == want.num_groups
    wk, wkv = want.keys[0]
    for g, (fr, key, s, c) in enumerate(merged):
        assert key[1] == bool(wkv[g])
        if key[1]:
            assert key[0] == int(wk[g]), "ranking by global first row restores the reference's group order"
        assert c == int(want.aggs[1][0][0][g])
        if c:
            assert s == want.aggs[0][0][0][g]


def test_choose_splitters_clamps_tail_keys():
    from arrow1_b200 import dist as bd
    big = torch.full((100,), (2**64 - 1) ^ (1 << 63) - 0, dtype=torch.int64)  # biased null keys
    spl = bd.choose_splitters(big, 4)
    assert (spl <= bd._TAIL_KEY_BIASED).all()
    assert bd.choose_splitters(big[:0], 3).numel() == 2
