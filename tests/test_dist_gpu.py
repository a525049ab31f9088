"""Multi-GPU forms on real GPUs (NCCL, one process per GPU): needs >= 2 devices, skipped otherwise.
The sample sort runs both ways — the range partition fused with the exchange over NVLink peer mappings
and the NCCL all-to-all form — and both must equal the oracle's sort of the concatenation."""
import os
import pickle
import tempfile

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
from arrow1_b200.columns import Column
from test_dist_cpu import U64, _free_port, _sort_cases, shard

pytestmark = pytest.mark.gpu
ORACLE = "reference" if oracle.have_reference() else "port"


def _worker(rank, world, port, job, path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda:%d" % rank))
    from arrow1_b200 import compute, dist as bd
    ops = bd.CudaOps(compute.default_context(rank))
    res = []
    for case in job:
        col, base = case["shards"][rank]
        bd.FUSED_EXCHANGE = case["fused"]
        if case.get("no_peer"):   # a node without peer mappings: every rank must agree to fall back to NCCL
            os.environ["B200_NO_PEER"] = "1"
            ops.ctx._peer_sort, ops.ctx._peer_sort_failed = None, False
        elif os.environ.pop("B200_NO_PEER", None):
            ops.ctx._peer_sort, ops.ctx._peer_sort_failed = None, False
        out, off = bd.dist_sort_indices(ops, col.to_device(), case["order"], base)
        res.append((out.cpu().numpy().view(U64).copy(), off))
    with open(path + ".%d" % rank, "wb") as f:
        pickle.dump(res, f)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_dist_sort_indices_fused_and_all_to_all_match_oracle(world):
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    rng = np.random.default_rng(300 + world)
    cases = _sort_cases(rng)
    n = 300_000   # many tiles per rank, uneven runs
    f = np.floor(rng.random(n) * 4000.0) - 2000.0
    f[rng.integers(0, n, 3000)] = np.nan
    cases.append(Column.from_numpy(f, rng.random(n) > 0.02))
    cases.append(Column.from_numpy(rng.integers(-2**63, 2**63 - 1, size=n, dtype=np.int64)))
    job, wants = [], []
    job.append({"order": 0, "fused": True, "no_peer": True, "shards": shard(cases[0], world)})
    wants.append(oracle.sort_indices(cases[0], 0, kind=ORACLE))
    for col in cases:
        for order in (0, 1):
            for fused in (True, False):
                job.append({"order": order, "fused": fused, "shards": shard(col, world)})
                wants.append(oracle.sort_indices(col, order, kind=ORACLE))
    port = _free_port()
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "out")
        mp.spawn(_worker, args=(world, port, job, path), nprocs=world, join=True)
        results = [pickle.load(open(path + ".%d" % r, "rb")) for r in range(world)]
    for ci, want in enumerate(wants):
        parts = [results[r][ci] for r in range(world)]
        assert [p[1] for p in parts] == list(np.cumsum([0] + [len(p[0]) for p in parts[:-1]])), "slice offsets"
        np.testing.assert_array_equal(np.concatenate([p[0] for p in parts]), want, err_msg="case %d" % ci)


# ---- grouped aggregates over NCCL: the low-cardinality merge and the hash-repartitioning all-to-all ---------------
def _host_groups(out):
    return ([[(c.to_host().values().copy(), c.to_host().valid().copy()) for c in a] for a in out.aggs],
            [(c.to_host().values().copy(), c.to_host().valid().copy()) for c in out.keys], out.num_groups)


def _gb_worker(rank, world, port, job, path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda:%d" % rank))
    from arrow1_b200 import compute, dist as bd
    ops = bd.CudaOps(compute.default_context(rank))
    res = []
    for case in job:
        (kcols, acols), base = case["shards"][rank]
        dk = [k.to_device("cuda:%d" % rank) for k in kcols]
        da = {}
        dargs = []
        for a in acols:   # aggregates sharing a host column share the device column
            if id(a) not in da:
                da[id(a)] = a.to_device("cuda:%d" % rank)
            dargs.append(da[id(a)])
        if case["form"] == "merge":
            out = bd.dist_group_by(ops, dk, case["aggs"], dargs, base, case["hint"])
            res.append(("merge", _host_groups(out)))
        else:
            sg = bd.dist_group_by_repartition(ops, dk, case["aggs"], dargs, base, case["hint"])
            res.append(("repartition", _host_groups(sg.out), sg.first_rows.cpu().numpy().view(U64)[: sg.num_groups].copy()))
    with open(path + ".%d" % rank, "wb") as f:
        pickle.dump(res, f)
    dist.barrier()
    dist.destroy_process_group()


def _shard_table(keys, args, world):
    """the same uneven row ranges for every column; aggregates that share a column keep sharing it"""
    ks = [shard(k, world) for k in keys]
    uniq = {}
    for a in args:
        if id(a) not in uniq:
            uniq[id(a)] = shard(a, world)
    out = []
    for r in range(world):
        out.append((([k[r][0] for k in ks], [uniq[id(a)][r][0] for a in args]), ks[0][r][1]))
    return out


@pytest.mark.parametrize("world", [2, 4])
def test_dist_group_by_merge_and_repartition_match_oracle(world):
    """dist_group_by (C3 shape: local tables, all-gather of the slots, merge in rank order) and
    dist_group_by_repartition (C4 shape: hash partition, NCCL all-to-all, disjoint key sets; also through the
    partition-then-aggregate consume) on real GPUs against the reference on the whole columns: group order, keys, counts,
    min/max bit for bit, float sums within 1e-12."""
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    from arrow1_b200 import columns as cols
    from arrow1_b200.columns import AggSpec
    from helpers import GroupRes, abs_sum_scales, assert_group_equal
    rng = np.random.default_rng(700 + world)
    cases = []
    n = 400_000
    k = Column.from_numpy(rng.integers(0, 1000, size=n).astype(np.int32))
    v = Column.from_numpy(rng.random(n) * 1e14, rng.random(n) > 0.01)
    aggs3 = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_COUNT), AggSpec(cols.HASH_MIN_MAX)]
    cases.append(("merge", [k], aggs3, [v, v, v], 1000))
    k2 = Column.from_numpy(rng.integers(0, 5, size=n).astype(np.uint8), rng.random(n) > 0.05)
    cases.append(("merge", [k2, k], [AggSpec(cols.HASH_MEAN), AggSpec(cols.HASH_COUNT, 1)], [v, v], 0))
    with np.errstate(over="ignore"):
        k64 = Column.from_numpy((rng.integers(0, 150_000, size=n).astype(np.uint64) * np.uint64(0x9E3779B97F4A7C15)).view(np.int64))
    aggs4 = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_MEAN)]
    cases.append(("repartition", [k64], aggs4, [v, v], 150_000))
    kn = Column.from_numpy(rng.integers(-40_000, 40_000, size=n).astype(np.int64), rng.random(n) > 0.03)
    iv = Column.from_numpy(rng.integers(-10**9, 10**9, size=n).astype(np.int64), rng.random(n) > 0.1)
    cases.append(("repartition", [kn], [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_MIN_MAX), AggSpec(cols.HASH_COUNT)], [iv, iv, iv], 0))
    job = [{"form": f, "aggs": a, "hint": h, "shards": _shard_table(ks, ar, world)} for f, ks, a, ar, h in cases]
    port = _free_port()
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "out")
        mp.spawn(_gb_worker, args=(world, port, job, path), nprocs=world, join=True)
        results = [pickle.load(open(path + ".%d" % r, "rb")) for r in range(world)]
    for ci, (form, ks, aggs, args, _) in enumerate(cases):
        want = oracle.groupby(ks, aggs, args, kind=ORACLE)
        b = GroupRes(want.num_groups, want.aggs, want.keys)
        scales = abs_sum_scales(ks, aggs, args, ORACLE)
        if form == "merge":
            for r in range(world):   # every rank holds the whole result
                ra, rk, g = results[r][ci][1]
                assert_group_equal(GroupRes(g, ra, rk), b, aggs, rtol=1e-12, scales=scales)
        else:
            parts = [results[r][ci] for r in range(world)]
            first = np.concatenate([p[2] for p in parts])
            order = np.argsort(first, kind="stable")
            assert len(np.unique(first)) == len(first)
            n_aggs_out = len(parts[0][1][0])
            ra = []
            for j in range(n_aggs_out):
                ra.append([(np.concatenate([p[1][0][j][q][0] for p in parts])[order], np.concatenate([p[1][0][j][q][1] for p in parts])[order])
                           for q in range(len(parts[0][1][0][j]))])
            rk = [(np.concatenate([p[1][1][q][0] for p in parts])[order], np.concatenate([p[1][1][q][1] for p in parts])[order])
                  for q in range(len(ks))]
            assert_group_equal(GroupRes(len(first), ra, rk), b, aggs, rtol=1e-12, scales=scales)
