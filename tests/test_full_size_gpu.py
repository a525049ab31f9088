"""BASELINE.json's configurations at their FULL sizes, checked through size-independent properties computed on
the device (torch is the checker here, nothing else): the oracle cannot run 10^8..10^9 rows in test time, so the
parity tests proper run at reduced sizes against the oracle and these close the gap at the quoted sizes."""
import numpy as np
import pytest
import torch

from arrow1_b200 import capi, columns as cols, synth
from arrow1_b200.columns import AggSpec

pytestmark = pytest.mark.gpu


def _bits(buf, n):
    """LSB-first bitmap (device uint8 tensor) -> bool tensor[n]"""
    w = torch.tensor([1, 2, 4, 8, 16, 32, 64, 128], dtype=torch.uint8, device=buf.device)
    return ((buf[: (n + 7) // 8].unsqueeze(1) & w) != 0).reshape(-1)[:n]


def _valid_bits(col, n):
    """validity bitmap of a device column -> bool tensor[n]"""
    if col.validity is None:
        return torch.ones(n, dtype=torch.bool, device=col.data.device)
    return _bits(col.validity, n)


@pytest.mark.parametrize("dtype", ["int64", "float64"])
def test_sort_config2_full_size(ctx, dtype):
    """C2: 100 M rows, 1 % nulls.  The result is a permutation; valid rows come first in non-decreasing key order
    with ties in increasing row order (stability); the nulls follow in increasing row order."""
    n = 100_000_000
    col = synth.Gen("cuda:0").sort_c2(n, dtype)
    idx = capi.sort_indices(ctx, col, 0)
    assert idx.numel() == n
    seen = torch.zeros(n, dtype=torch.bool, device=idx.device)
    seen[idx] = True
    assert bool(seen.all()), "not a permutation"
    del seen
    valid = _valid_bits(col, n)
    n_valid = int(valid.sum().item())
    assert n_valid == n - col.null_count
    taken_valid = valid[idx]
    assert bool(taken_valid[:n_valid].all()) and not bool(taken_valid[n_valid:].any()), "nulls are not last"
    vals = col.data[: n * 8].view(torch.int64 if dtype == "int64" else torch.float64)
    k = vals[idx[:n_valid]]
    i = idx[:n_valid]
    lt, eq = k[:-1] < k[1:], k[:-1] == k[1:]
    assert bool((lt | (eq & (i[:-1] < i[1:]))).all()), "not sorted / not stable"
    tail = idx[n_valid:]
    assert bool((tail[:-1] < tail[1:]).all()), "nulls not in row order"


def test_filter_take_config1_scaled_full_size(ctx):
    """C1's generator at 100 M rows: Filter(values, mask) == Take(values, positions of the set mask bits),
    validity, null count and every valid value — two independent kernels agreeing at full size."""
    n = 100_000_000
    values, mask = synth.Gen("cuda:0").filter_c1(n)
    out = capi.filter(ctx, values, mask, 0)
    sel = _bits(mask.data, n)   # the mask has no nulls: selected rows = set data bits
    pos = torch.nonzero(sel).reshape(-1)
    assert out.length == pos.numel()
    took = capi.take(ctx, values, pos, True)
    m = out.length
    v_f = _valid_bits(out.col, m) if out.has_validity else torch.ones(m, dtype=torch.bool, device=pos.device)
    assert torch.equal(v_f, _valid_bits(took, m))
    assert took.null_count == int((~v_f).sum().item())
    # under a null, Filter keeps the source bytes (vector_selection.cc:700-712) and Take writes zero (:316-320)
    a, b = out.col.data[: m * 8].view(torch.int64), took.data[: m * 8].view(torch.int64)
    assert torch.equal(a[v_f], b[v_f])
    assert bool((b[~v_f] == 0).all())


def test_groupby_config3_full_size(ctx):
    """C3: 1 B rows, 1 K int32 keys, float64 values, 1 % nulls; sum + count + min_max.  Checksums: 1000 groups in
    first-occurrence order of the key stream; counts add up to the valid rows; the sum of the group sums equals the
    column's sum within 1e-12; the min of mins / max of maxes are the column's; a 50 M-row prefix run through the same
    path gives the same group order."""
    n = 1_000_000_000
    key, val = synth.Gen("cuda:0").groupby_c3(n, 1000)
    aggs = [AggSpec(cols.HASH_SUM), AggSpec(cols.HASH_COUNT), AggSpec(cols.HASH_MIN_MAX)]
    got = capi.group_by(ctx, [key], aggs, [val] * 3, expected_groups=1000)
    assert got.num_groups == 1000
    keys = got.keys[0].values()
    assert sorted(keys.tolist()) == list(range(1000))
    valid = _valid_bits(val, n)
    v = val.data[: n * 8].view(torch.float64)
    counts = got.aggs[1][0].values()
    assert int(counts.sum()) == int(valid.sum().item()) == n - val.null_count
    sums = got.aggs[0][0].values()
    total = float(torch.where(valid, v, torch.zeros_like(v)).sum(dtype=torch.float64).item())
    assert abs(float(np.sum(sums)) - total) <= 1e-12 * total * 4   # two different summation orders of 10^9 terms
    mins, maxes = got.aggs[2][0].values(), got.aggs[2][1].values()
    big = torch.full_like(v, float("inf"))
    assert float(mins.min()) == float(torch.where(valid, v, big).min().item())
    assert float(maxes.max()) == float(torch.where(valid, v, -big).max().item())
    del big
    # group order = first occurrence: the first row of every key, ranked
    k32 = key.data[: n * 4].view(torch.int32)
    head = k32[:50_000_000].to(torch.int64)
    first = torch.full((1000,), n, dtype=torch.int64, device=head.device)
    first.scatter_reduce_(0, head, torch.arange(head.numel(), device=head.device), reduce="amin")
    assert int(first.max().item()) < head.numel(), "every key appears in the prefix"
    want_order = torch.argsort(first).cpu().numpy().astype(np.int32)
    np.testing.assert_array_equal(keys, want_order)
