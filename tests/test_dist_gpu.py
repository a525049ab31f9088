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
