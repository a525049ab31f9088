"""Seeded input generators shared by the CPU and GPU parity tests (shapes follow the reference's
randomised tests: vector_selection_test.cc:1618-1636, vector_sort_test.cc:550-624,
hash_aggregate_test.cc:704-723) plus the SURVEY §8d benchmark columns."""
import numpy as np

from arrow1_b200 import columns as cols
from arrow1_b200.columns import Column

ALL_NUMERIC = [cols.INT8, cols.UINT8, cols.INT16, cols.UINT16, cols.INT32, cols.UINT32, cols.INT64, cols.UINT64,
               cols.FLOAT, cols.DOUBLE]


def rand_values(rng, t, n, lo=None, hi=None, special=False):
    if t == cols.BOOL:
        return rng.random(n) < 0.5
    dt = np.dtype(cols.NP_OF[t])
    if dt.kind == "f":
        a = (rng.random(n) * 2e3 - 1e3).astype(dt) if lo is None else (rng.random(n) * (hi - lo) + lo).astype(dt)
        if special and n > 0:
            idx = rng.integers(0, n, size=max(n // 8, 1))
            pool = np.array([np.nan, -np.nan, 0.0, -0.0, np.inf, -np.inf, 1e-310, -1e-310, 1.0, 1.0], dtype=dt)
            a[idx] = pool[rng.integers(0, len(pool), size=len(idx))]
        return a
    info = np.iinfo(dt)
    lo_ = info.min if lo is None else lo
    hi_ = info.max if hi is None else hi
    a = rng.integers(lo_, hi_, size=n, dtype=dt, endpoint=True)
    if special and n > 0:
        idx = rng.integers(0, n, size=max(n // 8, 1))
        pool = np.array([info.min, info.max, 0, 1], dtype=dt)
        a[idx] = pool[rng.integers(0, len(pool), size=len(idx))]
    return a


def rand_column(rng, t, n, null_prob=0.0, offset=0, lo=None, hi=None, special=False, force_bitmap=False):
    """Column of n rows sliced at `offset` out of a parent of n+offset rows."""
    total = n + offset
    v = rand_values(rng, t, total, lo, hi, special)
    valid = None
    if null_prob > 0 or force_bitmap:
        valid = rng.random(total) >= null_prob
    return Column.from_numpy(v, valid, offset=offset, type=t)


def rand_mask(rng, n, true_prob, null_prob=0.0, offset=0):
    total = n + offset
    bits = rng.random(total) < true_prob
    valid = (rng.random(total) >= null_prob) if null_prob > 0 else None
    return Column.from_numpy(bits, valid, offset=offset)
