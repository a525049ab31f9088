"""compute.Pipeline: several host-buffer calls in flight (one b200_ctx + CUDA stream per lane) must return
exactly what the same calls return one at a time, and what the oracle returns."""
import numpy as np
import pytest

import oracle
from arrow1_b200 import columns as cols
from arrow1_b200.columns import Column
from gen import rand_column

pytestmark = pytest.mark.gpu
ORACLE = "reference" if oracle.have_reference() else "port"


def test_pipelined_sort_indices_match_oracle():
    from arrow1_b200 import compute
    rng = np.random.default_rng(2024)
    batches = []
    for i in range(7):
        t = (cols.INT64, cols.DOUBLE, cols.UINT32)[i % 3]
        batches.append(compute.pinned_column(rand_column(rng, t, 150_000 + 4097 * i, 0.1 * (i % 3), special=(i % 2 == 0))))
    pipe = compute.Pipeline(depth=3)
    try:
        for order in (compute.ASCENDING, compute.DESCENDING):
            pending = [pipe.submit(compute.SortIndices, b, order) for b in batches]
            for b, p in zip(batches, pending):
                got = p.result()
                assert got.dtype == np.uint64
                np.testing.assert_array_equal(got, oracle.sort_indices(b, order, 0, kind=ORACLE))
        assert pipe.launches > 0
        # copies started ahead of the call (Pipeline.stage): same results
        staged = [pipe.stage(b) for b in batches]
        pending = [pipe.submit(compute.SortIndices, st, compute.ASCENDING) for st in staged]
        for b, p in zip(batches, pending):
            np.testing.assert_array_equal(p.result(), oracle.sort_indices(b, compute.ASCENDING, 0, kind=ORACLE))
    finally:
        pipe.close()


def test_pipelined_filter_matches_oracle_and_serial_calls():
    from arrow1_b200 import compute
    rng = np.random.default_rng(77)
    pipe = compute.Pipeline(depth=2)
    try:
        work = []
        for i in range(6):
            n = 200_000 + 1001 * i
            v = compute.pinned_column(rand_column(rng, cols.INT64, n, 0.2 if i % 2 else 0.0))
            m = compute.pinned_column(Column.from_numpy(rng.random(n) < 0.4, (rng.random(n) > 0.1) if i % 3 == 0 else None))
            for sel in (compute.DROP, compute.EMIT_NULL):
                if i % 2:
                    work.append((v, m, sel, pipe.submit(compute.Filter, v, m, sel)))
                else:   # both columns staged on one lane ahead of the call
                    sv, sm = pipe.stage(v, m)
                    work.append((v, m, sel, pipe.submit(compute.Filter, sv, sm, sel)))
        for v, m, sel, p in work:
            got = p.result()
            want = oracle.filter(v, m, sel, kind=ORACLE)
            serial = compute.Filter(v, m, sel)
            assert got.length == want.length == serial.length
            assert got.null_count == want.null_count
            np.testing.assert_array_equal(got.values(), serial.values())
            valid = got.valid() if got.validity is not None else np.ones(got.length, dtype=np.bool_)
            np.testing.assert_array_equal(valid, want.valid())
            np.testing.assert_array_equal(got.values()[valid], want.values()[valid])
    finally:
        pipe.close()
