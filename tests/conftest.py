import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from arrow1_b200 import columns as cols  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")

TYPE_BY_NAME = {v: k for k, v in cols.NAMES.items()}


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


def _num(x):
    if isinstance(x, str):
        return float(x.replace("Inf", "inf").replace("NaN", "nan"))
    return x


def column_from_json(values, type_name, offset=0):
    """JSON list (null -> null, "NaN"/"Inf"/"-0.0" strings -> floats) -> host Column."""
    t = TYPE_BY_NAME[type_name]
    valid = np.array([v is not None for v in values], dtype=np.bool_)
    if t == cols.BOOL:
        arr = np.array([bool(v) if v is not None else False for v in values], dtype=np.bool_)
    else:
        arr = np.array([_num(v) if v is not None else 0 for v in values], dtype=cols.NP_OF[t])
    has_nulls = not valid.all()
    return cols.Column.from_numpy(arr, valid if has_nulls else None, offset=offset, type=t)


def expected_from_json(values, type_name):
    t = TYPE_BY_NAME[type_name]
    valid = np.array([v is not None for v in values], dtype=np.bool_)
    if t == cols.BOOL:
        arr = np.array([bool(v) if v is not None else False for v in values], dtype=np.bool_)
    else:
        arr = np.array([_num(v) if v is not None else 0 for v in values], dtype=cols.NP_OF[t])
    return arr, valid


@pytest.fixture(scope="session")
def ctx():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from arrow1_b200 import capi
    return capi.Context(0)


def splitmix64(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15)) & np.uint64(0xFFFFFFFFFFFFFFFF)
    z = x
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def rand_u64(seed, n):
    """SURVEY §8d generator r(s, i) = splitmix64(s * 0x9E3779B97F4A7C15 + i)."""
    with np.errstate(over="ignore"):
        i = np.arange(n, dtype=np.uint64)
        return splitmix64(np.uint64(seed) * np.uint64(0x9E3779B97F4A7C15) + i)
