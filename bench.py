#!/usr/bin/env python
"""bench.py — throughput of the Filter / SortIndices / grouped-aggregate hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl reference]

One JSON line on stdout (rank 0).  A "step" is one pass of the workload's operation over one
batch of synthetic columns (arrow1_b200/synth.py, counter-based).  Workloads:

    sort_i64    SortIndices ascending, 100 M int64 rows, 1 % nulls          (BASELINE configs[1], default)
    sort_f64    SortIndices ascending, 100 M float64 rows, 1 % nulls        (BASELINE configs[1])
    filter      Filter, 100 M int64 rows, 50 % mask, 1 % nulls              (configs[0] shape at the timed size)
    filter_10m  Filter, 10 M rows                                           (configs[0] as written)
    groupby_c3  hash_sum+hash_count+hash_min_max, (int32 key, float64), 1 K keys, 1 B rows   (configs[2])
    groupby_c4  hash_sum+hash_mean, int64 keys, 100 M distinct, 1 B rows                     (configs[3])

`value` = rows/s of the whole job with inputs resident in HBM; `e2e` = the same through
arrow1_b200.compute (the reference-facing API) from pinned HOST buffers, copies inside the timed
region.  `roofline` describes the workload's dominant kernel (per-launch CUDA events from the
library's timing hook, taken over the timed steps); `cpu_baseline` is the reference's own CPU
kernels (oracle/_ref) timed on this box's host cores on a bounded sample.  The default run adds a
`workloads` object with the device-resident numbers of the other configs.

`--impl reference` times the reference's CPU implementation (oracle/_ref, else the oracle port)
of the same workload on the host cores and prints the same line with "impl": "reference".
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "rows/s & input GB/s (filter, sort_indices, hash_sum) at 1-8 B200; % HBM peak"
L2_BYTES = 126 * 1024 * 1024


def log(*a):
    print(*a, file=sys.stderr, flush=True)


_JSON_FD = None


def claim_stdout():
    """The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version banner on
    rank 0), so fd 1 is pointed at stderr for the whole run and the JSON line goes to the original stdout."""
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    sys.stdout.flush()
    if _JSON_FD is None:
        os.write(1, data)
    else:
        os.write(_JSON_FD, data)


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------------
# workload definitions

AGG_C3 = ("hash_sum", "hash_count", "hash_min_max")
AGG_C4 = ("hash_sum", "hash_mean")


class Workload:
    name = ""
    default_rows = 0
    kernel = ""          # dominant kernel (name in the library's timing hook)
    dtype = ""
    description = ""

    def __init__(self, rows):
        self.n = rows

    # generated columns (host when gen.device is None)
    def make(self, gen, start=0, n=None):
        raise NotImplementedError

    def input_bytes(self):      # the reference benchmarks' convention
        raise NotImplementedError

    def alg_bytes_per_launch(self, launches_per_step):   # dominant kernel, per launch
        raise NotImplementedError

    def prepare(self, ctx, cols_):   # device buffers reused across steps
        return None

    def step(self, ctx, cols_, bufs):
        raise NotImplementedError

    def e2e_step(self, host_cols):
        raise NotImplementedError

    def cpu_step(self, oracle, kind, host_cols):
        raise NotImplementedError


class SortWL(Workload):
    kernel = "sort_pass"

    def __init__(self, rows, dtype):
        super().__init__(rows)
        self.dt = dtype
        self.name = "sort_i64" if dtype == "int64" else "sort_f64"
        self.dtype = "u64 radix keys (%s column)" % dtype
        self.description = "SortIndices ascending, %d %s rows, 1%% nulls (at end), stable" % (rows, dtype)

    def make(self, gen, start=0, n=None):
        return (gen.sort_c2(n or self.n, self.dt, start),)

    def input_bytes(self):
        return self.n * 8

    def alg_bytes_per_launch(self, launches_per_step):
        # 8 live digit passes of a 64-bit column (DESIGN.md "Sort"): the first reads 8.125 B/row and writes key + row
        # number (12); passes 1, 2 read and write 12; pass 3 reads 12 and writes the upper key word + row (8); passes
        # 4..6 read and write 8; the last reads 8 and writes the 8-byte index
        total = self.n * (8.125 + 12 + 2 * 24 + 12 + 8 + 3 * 16 + 8 + 8)
        if getattr(self, "high_word_first", False):
            # high word first (sort.cu HwState): 4 narrow passes over digits 4..7 — the first reads the column and
            # writes upper word + row (8), two read and write 8, the last reads 8 and writes the index (8) and the
            # sorted upper word (4); the passes over digits 0..3 are launched and exit at once
            total = self.n * (8.125 + 8 + 2 * 16 + 8 + 12)
            return total / 4
        return total / max(launches_per_step, 1)

    def live_launches_per_step(self, launches_per_step):
        return 4 if getattr(self, "high_word_first", False) else launches_per_step

    def after_timing(self, ctx, bufs):
        """Which plan the sort took (decided on the device from the digit histograms)."""
        import ctypes as C
        from arrow1_b200 import capi
        hw, fell_back = C.c_int(0), C.c_int(0)
        ctx.check(capi.lib().b200_sort_indices_plan(ctx._h, C.c_void_p(bufs["scratch"].data_ptr()), C.byref(hw), C.byref(fell_back)))
        self.high_word_first = bool(hw.value) and not fell_back.value
        self.plan = {"high_word_first": bool(hw.value), "fell_back_to_8_passes": bool(fell_back.value)}

    def prepare(self, ctx, cols_):
        import torch
        from arrow1_b200 import capi
        return {"out": torch.empty(self.n, dtype=torch.int64, device="cuda:%d" % ctx.device),
                "scratch": ctx.empty(capi.sort_scratch_bytes(ctx, self.n))}

    def step(self, ctx, cols_, bufs):
        from arrow1_b200 import capi
        return capi.sort_indices(ctx, cols_[0], 0, 0, bufs["out"], bufs["scratch"])

    def e2e_step(self, host_cols):
        from arrow1_b200 import compute
        return compute.SortIndices(host_cols[0], compute.ASCENDING)

    def e2e_submit(self, pipe, host_cols):
        from arrow1_b200 import compute
        return pipe.submit(compute.SortIndices, host_cols[0], compute.ASCENDING)

    def cpu_step(self, oracle, kind, host_cols):
        return oracle.sort_indices(host_cols[0], 0, 0, kind=kind)


class SortMultiWL(Workload):
    """SortIndices(RecordBatch, [key ascending, value descending]) — two sort keys."""
    kernel = "sort_pass"
    name = "sort_2keys"
    dtype = "u64 radix keys (int32 and int64 columns)"

    def __init__(self, rows):
        super().__init__(rows)
        self.description = ("SortIndices by (int32 key in [0, 1000) ascending, int64 value descending, 1%% nulls last), "
                            "%d rows, stable" % rows)

    def make(self, gen, start=0, n=None):
        return (gen.groupby_c3(n or self.n, 1000, start)[0], gen.sort_c2(n or self.n, "int64", start))

    def input_bytes(self):
        return self.n * 12

    def alg_bytes_per_launch(self, launches_per_step):
        # value key: 8 live passes of (8 B key, 4 B row) pairs + 1 class pass; int32 key: 2 live passes
        # (values < 1000); every pass reads and writes 12 B/row, the last of a sort writes 8 B rows
        total = self.n * (11 * 24)
        if self._high_word_first():
            # the value key's sort took the upper-key-word-first plan (4 narrow passes, 68 B/row, see SortWL); the int32
            # key's pair sort: 2 live passes of 24 B/row
            return self.n * (68.125 + 2 * 24) / 6
        return total / max(launches_per_step, 1)

    def _high_word_first(self):
        # the tie repair only does work under that plan (0.27 ms per 100 M rows against 0.01 ms idle)
        kt, steps = getattr(self, "_kt", None), getattr(self, "_kt_steps", 1)
        return bool(kt) and kt.get("sort_tie_repair", (0.0, 0))[0] / max(steps, 1) > 0.001 * self.n / 1e6

    def live_launches_per_step(self, launches_per_step):
        return 6 if self._high_word_first() else launches_per_step

    def prepare(self, ctx, cols_):
        return {}

    def step(self, ctx, cols_, bufs):
        from arrow1_b200 import capi
        return capi.sort_indices_multi(ctx, list(cols_), [0, 1])

    def e2e_step(self, host_cols):
        from arrow1_b200 import compute
        return compute.SortIndices(list(host_cols), [compute.ASCENDING, compute.DESCENDING])

    def cpu_step(self, oracle, kind, host_cols):
        return oracle.sort_indices_multi(list(host_cols), [0, 1], kind=kind)


class TakeWL(Workload):
    """the gather that follows SortIndices — Take(values, SortIndices(values)), i.e. the sorted
    column itself; the indices are a random permutation of the rows."""
    kernel = "take"
    name = "take_sorted"
    dtype = "u64 (bit copy of int64) + bitmaps, uint64 indices"

    def __init__(self, rows):
        super().__init__(rows)
        self.description = "Take(values, SortIndices(values)): %d int64 rows gathered by a random permutation, 1%% nulls" % rows

    def make(self, gen, start=0, n=None):
        return (gen.sort_c2(n or self.n, "int64", start),)

    def input_bytes(self):
        return self.n * 16

    def alg_bytes_per_launch(self, launches_per_step):
        return self.n * (8 + 8 + 8 + 0.25)   # index + value in (plus its validity bit), value + validity bit out

    def prepare(self, ctx, cols_):
        from arrow1_b200 import capi
        return {"indices": capi.sort_indices(ctx, cols_[0], 0)}

    def step(self, ctx, cols_, bufs):
        from arrow1_b200 import capi
        return capi.take(ctx, cols_[0], bufs["indices"], True, want_null_count=True)

    def e2e_step(self, host_cols):
        from arrow1_b200 import compute
        from arrow1_b200.columns import Column
        idx = getattr(self, "_e2e_idx", None)
        if idx is None or idx[0] is not host_cols[0]:   # the permutation is an input of Take: made once, untimed
            perm = compute.SortIndices(host_cols[0], compute.ASCENDING)
            self._e2e_idx = idx = (host_cols[0], compute.pinned_column(Column.from_numpy(perm)))
        return compute.Take(host_cols[0], idx[1], True)

    def cpu_step(self, oracle, kind, host_cols):
        from arrow1_b200.columns import Column
        cache = self.__dict__.setdefault("_cpu_idx", {})   # one permutation per shard, made once (an input of Take)
        idx = cache.get(id(host_cols[0]))
        if idx is None:
            idx = cache[id(host_cols[0])] = (host_cols[0], Column.from_numpy(oracle.sort_indices(host_cols[0], 0, 0, kind=kind)))
        return oracle.take(host_cols[0], idx[1], True, kind=kind)


class FilterWL(Workload):
    kernel = "filter_apply"
    dtype = "u64 (bit copy of int64) + bitmaps"

    def __init__(self, rows, name):
        super().__init__(rows)
        self.name = name
        self.description = "Filter, %d int64 rows, 50%% boolean mask, 1%% value nulls, DROP" % rows
        self.out_len = None

    def make(self, gen, start=0, n=None):
        return gen.filter_c1(n or self.n, start)

    def input_bytes(self):
        return self.n * 8   # values buffer only (vector_selection_benchmark.cc:92-97)

    def alg_bytes_per_launch(self, launches_per_step):
        sel = (self.out_len or self.n // 2)
        return self.n * (8 + 1 / 8 + 1 / 8) + sel * (8 + 1 / 8)

    def prepare(self, ctx, cols_):
        from arrow1_b200 import capi
        return {"data": ctx.empty(self.n * 8), "validity": ctx.empty((self.n + 7) // 8 + 8),
                "scratch": capi.FusedFilterScratch(ctx, self.n)}

    def step(self, ctx, cols_, bufs):
        # reused buffers sized by the mask length; plan + apply back to back, the output length stays on the device
        # (b200_filter_fused; the form with the host round trip between them is what compute.Filter / e2e use)
        from arrow1_b200 import capi
        out = capi.filter_fused(ctx, cols_[0], cols_[1], 0, bufs["data"], bufs["validity"], self.n, bufs["scratch"], sync=False)
        return out

    def e2e_step(self, host_cols):
        from arrow1_b200 import compute
        return compute.Filter(host_cols[0], host_cols[1], compute.DROP)

    def e2e_submit(self, pipe, host_cols):
        from arrow1_b200 import compute
        return pipe.submit(compute.Filter, host_cols[0], host_cols[1], compute.DROP)

    def cpu_step(self, oracle, kind, host_cols):
        return oracle.filter(host_cols[0], host_cols[1], 0, kind=kind)


class GroupByWL(Workload):
    kernel = "groupby_consume"

    def __init__(self, rows, which):
        super().__init__(rows)
        self.which = which
        self.name = "groupby_" + which
        if which == "c3":
            self.n_keys, self.agg_names = 1000, AGG_C3
            self.dtype = "f64 sums / u64 counts, int32 keys"
            self.description = "hash_sum+hash_count+hash_min_max over %d rows of (int32 key, float64 value), 1K keys, 1%% nulls" % rows
        else:
            self.n_keys, self.agg_names = min(100_000_000, max(rows // 10, 1)), AGG_C4
            self.dtype = "f64 sums / u64 counts, int64 keys"
            self.description = "hash_sum+hash_mean over %d rows of (int64 key, float64 value), %d distinct keys" % (rows, self.n_keys)

    def _specs(self):
        from arrow1_b200 import columns as cols
        kind = {"hash_sum": cols.HASH_SUM, "hash_count": cols.HASH_COUNT, "hash_min_max": cols.HASH_MIN_MAX,
                "hash_mean": cols.HASH_MEAN}
        return [cols.AggSpec(kind[a]) for a in self.agg_names]

    def make(self, gen, start=0, n=None):
        if self.which == "c3":
            return gen.groupby_c3(n or self.n, self.n_keys, start)
        return gen.groupby_c4(n or self.n, self.n_keys, start)

    def input_bytes(self):
        return self.n * (12 if self.which == "c3" else 16)

    def alg_bytes_per_launch(self, launches_per_step):
        per_row = 4 + 8 + 1 / 8 if self.which == "c3" else 16
        return self.n * per_row / max(launches_per_step, 1)

    def step(self, ctx, cols_, bufs):
        from arrow1_b200 import capi
        k, v = cols_
        opts = None
        if os.environ.get("B200_BENCH_GB_OPTIONS"):   # experiments: "name=value,name=value" -> b200_groupby_set_option
            opts = {a.split("=")[0]: int(a.split("=")[1]) for a in os.environ["B200_BENCH_GB_OPTIONS"].split(",")}
        return capi.group_by(ctx, [k], self._specs(), [v] * len(self.agg_names), expected_groups=self.n_keys, options=opts)

    def e2e_step(self, host_cols):
        from arrow1_b200 import compute
        k, v = host_cols
        return compute.GroupBy([v] * len(self.agg_names), [k], self._specs(), expected_groups=self.n_keys)

    def cpu_step(self, oracle, kind, host_cols):
        k, v = host_cols
        return oracle.groupby([k], self._specs(), [v] * len(self.agg_names), kind=kind, max_groups=min(k.length, self.n_keys + 2))


class Q1WL(Workload):
    """Config C5: TPC-H Q1-shaped pipeline on synthetic lineitem, UNFUSED through the public operations:
    less_equal(shipdate) -> Filter of the 5 used columns (one shared plan) -> GroupBy(returnflag, linestatus)."""
    kernel = "groupby_consume"
    name = "q1"
    dtype = "f64 sums / means, u8 keys, date32 predicate"

    def __init__(self, rows):
        super().__init__(rows)
        self.description = ("Q1-shaped pipeline over %d synthetic lineitem rows: shipdate <= 1998-09-02 (~96%% selective), Filter of "
                            "returnflag/linestatus/quantity/extendedprice/discount, GroupBy(returnflag, linestatus) "
                            "sum x2, mean x3, count; unfused (each operation reads and writes HBM)" % rows)

    def make(self, gen, start=0, n=None):
        return gen.q1_lineitem(n or self.n, start)

    def input_bytes(self):
        return self.n * 30          # 4 (date) + 2 (keys) + 3 x 8 (measures)

    def alg_bytes_per_launch(self, launches_per_step):
        # group-by consume reads the two filtered key columns and the three filtered measures once
        return 0.964 * self.n * 26 / max(launches_per_step, 1)

    def step(self, ctx, cols_, bufs):
        from arrow1_b200 import compute
        ship, rf, ls, qty, price, disc = cols_
        if self.fused == "one_pass":
            return compute.Q1PipelineFused(ship, rf, ls, [qty, price, disc], ctx)
        if self.fused:
            return compute.Q1PipelineSelected(ship, rf, ls, [qty, price, disc], ctx)
        return compute.Q1Pipeline(ship, rf, ls, [qty, price, disc], ctx)

    fused = False

    def e2e_step(self, host_cols):
        from arrow1_b200 import compute
        ctx = compute.default_context()
        dev, nbytes = [], 0
        for c in host_cols:
            d, b = compute._h2d(c, ctx)
            dev.append(d)
            nbytes += b
        fn = {False: compute.Q1Pipeline, True: compute.Q1PipelineSelected, "one_pass": compute.Q1PipelineFused}[self.fused]
        out = fn(dev[0], dev[1], dev[2], dev[3:], ctx)
        d2h = 0
        for cols__ in out.aggs + [out.keys]:
            for c in cols__:
                d2h += c.to_host().nbytes()
        compute.Transfer.h2d, compute.Transfer.d2h = nbytes, d2h
        return out

    def cpu_step(self, oracle, kind, host_cols):
        import numpy as np
        from arrow1_b200 import compute
        from arrow1_b200.columns import Column
        ship, rf, ls, qty, price, disc = host_cols
        mask = Column.from_numpy(ship.values() <= compute.Q1_SHIPDATE_CUTOFF)
        f = [Column.from_numpy(oracle.filter(c, mask, 0, kind=kind).values().copy()) for c in (rf, ls, qty, price, disc)]
        f.append(Column.from_numpy(f[3].values() * (1.0 - f[4].values())))   # disc_price: two IEEE operations, like multiply / subtract
        specs = compute.q1_aggregates()
        return oracle.groupby([f[0], f[1]], [a for _, a in specs], [f[2 + i] for i, _ in specs], kind=kind, max_groups=8)


def make_workload(name, rows=None):
    if name == "q1":
        return Q1WL(rows or 600_000_000)
    if name == "q1_selected":
        w = Q1WL(rows or 600_000_000)
        w.fused, w.name = True, "q1_selected"
        w.description += "; Filter folded into the aggregation (b200_groupby_consume_selected)"
        return w
    if name == "q1_fused":
        w = Q1WL(rows or 600_000_000)
        w.fused, w.name, w.kernel = "one_pass", "q1_fused", "groupby_fused"
        w.description = w.description.replace("; unfused (each operation reads and writes HBM)",
                                              "; predicate, price*(1-discount) and the 7 aggregates in ONE pass over the 6 input columns "
                                              "(b200_groupby_consume_fused, 30 B/row)")
        w.alg_bytes_per_launch = lambda launches_per_step, n=w.n: n * 30 / max(launches_per_step, 1)
        return w
    if name == "sort_i64":
        return SortWL(rows or 100_000_000, "int64")
    if name == "sort_f64":
        return SortWL(rows or 100_000_000, "float64")
    if name == "take_sorted":
        return TakeWL(rows or 100_000_000)
    if name == "sort_2keys":
        return SortMultiWL(rows or 100_000_000)
    if name == "filter":
        return FilterWL(rows or 100_000_000, "filter")
    if name == "filter_10m":
        return FilterWL(rows or 10_000_000, "filter_10m")
    if name == "groupby_c3":
        return GroupByWL(rows or 1_000_000_000, "c3")
    if name == "groupby_c4":
        return GroupByWL(rows or 1_000_000_000, "c4")
    raise SystemExit("unknown workload %s" % name)


# ------------------------------------------------------------------------------------------------
# clocks

class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.path = index, None, None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return None
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 7:
                    continue
                try:
                    sm.append(float(f[0]))
                    mx.append(float(f[1]))
                except ValueError:
                    continue
                for nm, val in zip(names, f[3:7]):
                    if val.lower().startswith("active"):
                        reasons.add(nm)
        finally:
            try:
                os.unlink(self.path)
            except OSError:
                pass
        if not sm:
            return None
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# reference / cpu baseline

def host_threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


# rows up to which the CPU arm processes the WHOLE configuration (P shards of rows / P each); above it a bounded sample
CPU_FULL_ROWS = 200_000_000
CPU_SAMPLE_PER_THREAD = {"groupby_c3": 8_000_000, "groupby_c4": 4_000_000, "q1": 4_000_000, "q1_selected": 4_000_000,
                         "q1_fused": 4_000_000}


def cpu_shard_rows(wl, threads):
    """(rows per shard, whole configuration?)  SURVEY §8(d): P shards of the SAME rows, rows / P each."""
    if wl.n <= CPU_FULL_ROWS:
        return (wl.n + threads - 1) // threads, True
    return min(CPU_SAMPLE_PER_THREAD.get(wl.name, 4_000_000), wl.n), False


def run_cpu(wl, steps, warmup, threads):
    """Times the reference's CPU kernels on `threads` row-range shards (the kernels are single-threaded by
    construction, compute/exec.h:86-92), one reference call per shard and step on its own host thread:
    rows/s = rows of all shards / time.  Up to CPU_FULL_ROWS the shards cover the whole configuration (100 M rows
    for the sort: P shards of 100 M / P rows); the 600 M - 1 B row configurations are sampled.  Sort and group-by
    shards are not merged, i.e. an upper bound for the CPU."""
    import oracle
    from arrow1_b200 import synth
    kind = "reference" if oracle.have_reference() else "port"
    if kind == "port":
        oracle.build()
    per, full = cpu_shard_rows(wl, threads)
    gen = synth.Gen(None)
    bounds = [(i * per, min(per, wl.n - i * per)) for i in range(threads) if i * per < wl.n] if full else \
             [(i * per, per) for i in range(threads)]
    shards = [wl.make(gen, start=lo, n=ln) for lo, ln in bounds]
    subs = {ln: make_workload(wl.name, ln) for _, ln in bounds}
    for sub in subs.values():   # shards of the SAME columns: keep the parent's key cardinality
        if hasattr(wl, "n_keys"):
            sub.n_keys = wl.n_keys
    pool = ThreadPoolExecutor(max_workers=threads)

    def one():
        t0 = time.perf_counter()
        list(pool.map(lambda x: subs[x[0][1]].cpu_step(oracle, kind, x[1]), zip(bounds, shards)))
        return time.perf_counter() - t0

    for _ in range(warmup):
        one()
    times = [one() for _ in range(steps)]
    pool.shutdown()
    total = sum(times)
    rows = sum(ln for _, ln in bounds)
    what = ("the whole configuration: %d rows as %d row-range shards of %d rows" % (rows, len(bounds), per)) if full else \
           ("a bounded sample: %d shards x %d rows of the %d-row configuration" % (len(bounds), per, wl.n))
    return {"rows_per_s": rows * steps / total, "ms_per_step": 1e3 * total / steps, "kind": kind, "cores": threads,
            "rows": rows, "full_config": full,
            "sample": "%s (%s columns, same generator), one reference call per shard per step on its own host thread, "
                      "%d steps after %d warm-up; shards not merged" % (what, wl.name, steps, warmup)}


def workload_config(wl, world):
    """`config` of the JSON line — the same object on the GPU arm and on the reference arm."""
    return {"workload": wl.name, "description": wl.description, "rows_per_gpu": wl.n, "rows_total": wl.n * world,
            "l2": "inputs (%.0f MB per step) larger than the 126 MB L2; no explicit flush" % (wl.input_bytes() / 1e6)
            if wl.input_bytes() > 2 * L2_BYTES else "inputs fit in L2: numbers are L2-warm",
            "parallelism": "row-range shards, one process per GPU" if world > 1 else "single GPU"}


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    wl = make_workload(args.workload, args.rows)
    threads = host_threads()
    r = run_cpu(wl, max(args.steps, 1), max(args.warmup, 0), threads)
    line = {
        "metric": METRIC, "impl": "reference", "value": r["rows_per_s"], "unit": "rows/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": wl.dtype, "data": "synthetic",
        "config": workload_config(wl, world),
        "cpu_baseline": {"value": r["rows_per_s"], "unit": "rows/s", "cores": r["cores"], "kind": r["kind"],
                         "sample": r["sample"], "rows_per_step": r["rows"], "whole_configuration": r["full_config"]},
        "e2e": {"value": r["rows_per_s"], "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "input_gbs": r["rows_per_s"] * wl.input_bytes() / wl.n / 1e9,
        "gpu_launches": 0,
    }
    emit(line)
    return 0


# ------------------------------------------------------------------------------------------------
# GPU arm

def kernel_times(ctx, names):
    import ctypes as C
    from arrow1_b200 import capi
    out = {}
    for nm in names:
        ms, n = C.c_double(0), C.c_int64(0)
        capi.lib().b200_ctx_kernel_time(ctx._h, nm.encode(), C.byref(ms), C.byref(n))
        out[nm] = (ms.value, n.value)
    return out


ALL_KERNELS = ("filter_plan", "filter_apply", "sort_hist", "sort_pass", "sort_tail", "sort_tie_repair", "sort_pass_round2", "groupby_consume", "compare_scalar", "arithmetic", "take", "partition_plan", "partition_scatter", "sort_partition",
               "groupby_partition", "groupby_import", "groupby_fused", "filter_fused", "groupby_wide_keys")


def time_device(ctx, wl, cols_, steps, warmup, world, dist_step=None):
    """W warm-up + K timed steps (CUDA events on the launching stream, max over ranks when world > 1, the library's
    per-launch event hook OFF so that launch-bound steps are not charged two event records per kernel), then a
    second pass of up to 5 steps with the hook on for the per-kernel times.  Returns per-step ms, kernel times
    {name: (total ms, launches)} of the second pass, its step count, launches per timed region, buffers."""
    import torch
    from arrow1_b200 import capi
    bufs = wl.prepare(ctx, cols_)
    fn = (lambda: dist_step(ctx, wl, cols_, bufs)) if dist_step else (lambda: wl.step(ctx, cols_, bufs))
    for _ in range(warmup):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()
    l0 = ctx.launches
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = e0.elapsed_time(e1)
    launches = ctx.launches - l0
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    k2 = max(1, min(steps, 5))
    capi.lib().b200_ctx_set_kernel_timing(ctx._h, 1)
    for _ in range(k2):
        fn()
    torch.cuda.synchronize()
    kt = kernel_times(ctx, ALL_KERNELS)
    capi.lib().b200_ctx_set_kernel_timing(ctx._h, 0)
    if hasattr(wl, "after_timing") and dist_step is None:
        wl.after_timing(ctx, bufs)
    if world > 1:
        dist.barrier()
    return ms / steps, kt, k2, launches, bufs


def roofline_of(wl, kt, steps, peak, peak_src):
    total_ms, n = kt.get(wl.kernel, (0.0, 0))
    if n == 0:
        return None
    wl._kt, wl._kt_steps = kt, steps
    live = wl.live_launches_per_step(n / steps) if hasattr(wl, "live_launches_per_step") else n / steps
    per_launch_ms = total_ms / steps / max(live, 1e-9)
    alg = wl.alg_bytes_per_launch(n / steps)
    achieved = alg / (per_launch_ms * 1e-3) / 1e9
    return {"bound": "hbm", "kernel": wl.kernel, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "frac_of_8TBs": achieved / 8000.0, "traffic": None, "peak_source": peak_src,
            "launches_per_step": live, "avg_launch_ms": per_launch_ms, "algorithmic_bytes_per_launch": alg,
            "kernel_share_of_step_events": None}


def measure_e2e(wl, ctx, rank, world, local, steps, pipelined=True):
    """The same metric end to end through arrow1_b200.compute (the reference-facing API): pinned HOST columns in,
    host result out, the copies inside the timed region.  Serial (one call at a time) and, where the workload has
    an asynchronous form, with 3 calls in flight (compute.Pipeline)."""
    import torch
    from arrow1_b200 import compute, synth
    if world > 1:
        import torch.distributed as dist
        from arrow1_b200 import dist as b200_dist
    e2e_rows = wl.n if wl.n <= 200_000_000 else 200_000_000
    sub = make_workload(wl.name, e2e_rows)
    hgen = synth.Gen(None)
    t0 = time.time()
    host_cols = tuple(compute.pinned_column(c) for c in sub.make(hgen, start=rank * wl.n, n=e2e_rows))
    log("[rank %d] host columns for e2e (%d rows) in %.1f s" % (rank, e2e_rows, time.time() - t0))
    if world > 1:
        from arrow1_b200 import dist as b200_dist
        e2e_fn = lambda: b200_dist.bench_e2e_step(ctx, sub, host_cols)  # noqa: E731
    else:
        e2e_fn = lambda: sub.e2e_step(host_cols)  # noqa: E731
    for _ in range(2):
        r = e2e_fn()
        del r
    e2e_steps = max(3, min(steps, 10))

    def timed(loop, k):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        loop(k)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / k
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    def serial_loop(k):
        for _ in range(k):
            r = e2e_fn()
            del r

    serial_ms = timed(serial_loop, e2e_steps)
    h2d_b, d2h_b = compute.Transfer.h2d, compute.Transfer.d2h
    ems, depth = serial_ms, 1
    dist_sort = world > 1 and sub.name.startswith("sort")
    if pipelined and ((world == 1 and hasattr(sub, "e2e_submit")) or dist_sort):
        # the same calls, `depth` in flight (compute.Pipeline: one b200_ctx + stream per lane): every step
        # still copies its own inputs in and its own result out inside the timed region, but step i+1's
        # H2D, step i's kernels and step i-1's D2H overlap
        depth = 3
        pipe = compute.Pipeline(depth=depth, device=local)

        def piped_loop(k):
            pend = []
            # the copy-in of step i+1 is started before step i is submitted: a step blocks the submitting
            # thread wherever it needs a host-visible number (filter's output length, the sort's exchange
            # counts), and the next copy should already be running by then
            stage = lambda: tuple(pipe.stage(*host_cols)) if len(host_cols) > 1 else (pipe.stage(*host_cols),)  # noqa: E731
            nxt = stage()
            for i in range(k):
                cur = nxt
                if i + 1 < k:
                    nxt = stage()
                if dist_sort:   # same submission order on every rank: the lanes' collectives line up
                    pend.append(pipe.submit(b200_dist.SortIndicesSharded, cur[0], 0, rank * wl.n))
                else:
                    pend.append(sub.e2e_submit(pipe, cur))
                if len(pend) >= depth:
                    pend.pop(0).result()
            while pend:
                pend.pop(0).result()
            pipe.join()

        piped_loop(2 * depth)   # warm the lanes' allocator pools and pinned result buffers
        e2e_steps *= 2              # the fill and drain of the pipeline are inside the timed region
        ems = timed(piped_loop, e2e_steps)
        pipe.close()
    e2e = {"value": e2e_rows * world / (ems * 1e-3), "unit": "rows/s", "h2d_bytes_per_step": h2d_b,
           "d2h_bytes_per_step": d2h_b, "ms_per_step": ems, "rows_per_step_per_gpu": e2e_rows,
           "steps": e2e_steps, "calls_in_flight": depth, "serial_ms_per_step": serial_ms,
           "serial_value": e2e_rows * world / (serial_ms * 1e-3),
           "api": "arrow1_b200.compute (mirror of compute::Filter / SortIndices / internal::GroupBy)"
                  + ("; compute.Pipeline keeps %d calls in flight, one lane (b200_ctx + stream) each" % depth if depth > 1 else "")}
    del host_cols
    return e2e


def glue_timing(name, rows, steps=5):
    """The same workload through the reference's own API — compute::CallFunction / cuda::GroupBy over the CUDA
    FunctionRegistry of arrow_glue (C++), device-resident and host-to-host — timed by arrow_glue/bench_glue, a binary
    built where the reference's headers are and shipped prebuilt.  None when the binary or the workload is missing."""
    exe = os.path.join(ROOT, "arrow_glue", "bench_glue")
    if not os.path.exists(exe) or name not in ("sort_i64", "filter", "groupby_c3"):
        return None
    try:
        r = subprocess.run([exe, name, str(rows), str(steps)], capture_output=True, text=True, timeout=600)
        if r.returncode != 0:
            return {"error": (r.stderr or r.stdout)[-300:]}
        return json.loads(r.stdout.strip().splitlines()[-1])
    except Exception as ex:
        return {"error": repr(ex)[:300]}


def gpu_arm(args):
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback (use --impl reference for the CPU arm)")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        log("bench: --gpus %d but WORLD_SIZE %d; using WORLD_SIZE" % (args.gpus, world))
    torch.cuda.set_device(local)
    dist_step = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda:%d" % local))
        from arrow1_b200 import dist as b200_dist
        dist_step = b200_dist.bench_step
    from arrow1_b200 import capi, compute, synth
    ctx = compute.default_context(local)
    peak, peak_src = measured_peak()
    wl = make_workload(args.workload, args.rows)
    gen = synth.Gen("cuda:%d" % local)
    t0 = time.time()
    cols_ = wl.make(gen, start=rank * wl.n)
    torch.cuda.synchronize()
    log("[rank %d] generated %s in %.1f s" % (rank, wl.name, time.time() - t0))

    sampler = ClockSampler(local)
    sampler.start()
    ms_step, kt, kt_steps, launches, bufs = time_device(ctx, wl, cols_, args.steps, args.warmup, world, dist_step)
    clocks = sampler.stop()
    rows_total = wl.n * world
    value = rows_total / (ms_step * 1e-3)
    roof = roofline_of(wl, kt, kt_steps, peak, peak_src)
    if roof:
        roof["kernel_share_of_step_events"] = (kt[wl.kernel][0] / kt_steps) / ms_step
    traffic_file = os.path.join(ROOT, "profiles", "traffic.json")
    if roof and os.path.exists(traffic_file):
        try:
            key = wl.name + ("_high_word_first" if getattr(wl, "high_word_first", False) else "")
            roof["traffic"] = json.load(open(traffic_file)).get(key, {}).get("dram_bytes_per_launch")
        except Exception:
            pass

    # ---- e2e: host (pinned) -> device -> op -> host, through arrow1_b200.compute
    e2e = None
    if not args.no_e2e:
        e2e = measure_e2e(wl, ctx, rank, world, local, args.steps)

    # ---- multi-GPU: the step's result checked against the generator, once, outside the timed region
    self_check = None
    if world > 1:
        try:
            self_check = b200_dist.bench_self_check(ctx, wl, cols_, bufs)
        except Exception as ex:
            self_check = {"error": repr(ex)[:300]}

    # ---- the other configs, device-resident numbers only
    extras = {}
    if world > 1 and not args.no_extras:
        # the other three §8(e) rows at this world size: Filter (no collective), low-cardinality merge (all-gather of
        # the slots), high-cardinality hash repartition (NCCL all-to-all).  C4 is the configuration as written: 1 B rows
        # and 100 M distinct keys IN TOTAL, i.e. strong scaling; the others keep the rows per GPU.
        del cols_, bufs
        torch.cuda.empty_cache()
        collective = {"filter": "none", "groupby_c3": "all_gather of the occupied slots + merge in rank order",
                      "groupby_c4": "all_to_all_single per column + row numbers (hash repartition)"}
        for name in ("sort_i64", "filter", "groupby_c3", "groupby_c4"):
            if name == wl.name:
                continue
            try:
                rows2 = 1_000_000_000 // world if name == "groupby_c4" else None
                w2 = make_workload(name, rows2)
                if name == "groupby_c4":
                    w2.n_keys = 100_000_000
                    w2.description = "hash_sum+hash_mean over %d rows in total of (int64 key, float64 value), 100000000 distinct keys" % (w2.n * world)
                c2 = w2.make(gen, start=rank * w2.n)
                k2 = max(3, min(args.steps, 5))
                ms2, kt2, ks2, l2, b2 = time_device(ctx, w2, c2, k2, 3, world, dist_step)
                entry = None
                chk = b200_dist.bench_self_check(ctx, w2, c2, b2)
                if rank == 0:
                    r2 = roofline_of(w2, kt2, ks2, peak, peak_src)
                    entry = {"rows_per_gpu": w2.n, "rows_total": w2.n * world, "ms_per_step": ms2,
                             "rows_per_s": w2.n * world / (ms2 * 1e-3), "scaling": "strong" if name == "groupby_c4" else "weak",
                             "collective": collective.get(name, "all_gather (sample, counts) + peer stores over NVLink"),
                             "gpu_launches": l2, "roofline": r2, "description": w2.description, "self_check": chk,
                             "kernel_ms_per_step": {k: v[0] / ks2 for k, v in kt2.items() if v[1]}}
                    if name == "groupby_c4":
                        # bytes this rank sends in the all-to-all: (key 8 + value 8 + row number 8) of (world-1)/world of its rows
                        sent = w2.n * 24 * (world - 1) / world
                        entry["all_to_all_bytes_sent_per_gpu"] = sent
                extras[name] = entry
                del c2, b2
                torch.cuda.empty_cache()
            except Exception as ex:
                extras[name] = {"error": repr(ex)[:300]}
        cols_, bufs = None, None
    if rank == 0 and world == 1 and not args.no_extras:
        del cols_, bufs
        torch.cuda.empty_cache()
        for name in ("filter", "filter_10m", "sort_i64", "sort_f64", "sort_2keys", "take_sorted", "groupby_c3", "groupby_c4", "q1", "q1_selected", "q1_fused"):
            if name == wl.name:
                continue
            try:
                w2 = make_workload(name, None)
                c2 = w2.make(gen)
                k2 = max(3, min(args.steps, 10))
                ms2, kt2, ks2, l2, b2 = time_device(ctx, w2, c2, k2, 3, 1)
                r2 = roofline_of(w2, kt2, ks2, peak, peak_src)
                if r2:
                    r2["kernel_share_of_step_events"] = (kt2[w2.kernel][0] / ks2) / ms2
                    try:
                        r2["traffic"] = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(name + ("_high_word_first" if getattr(w2, "high_word_first", False) else ""), {}).get("dram_bytes_per_launch")
                    except Exception:
                        pass
                extras[name] = {"rows": w2.n, "ms_per_step": ms2, "rows_per_s": w2.n / (ms2 * 1e-3),
                                "input_gbs": w2.input_bytes() / (ms2 * 1e-3) / 1e9, "gpu_launches": l2,
                                "roofline": r2, "description": w2.description,
                                "kernel_ms_per_step": {k: v[0] / ks2 for k, v in kt2.items() if v[1]}}
                if hasattr(w2, "plan"):
                    extras[name]["plan"] = w2.plan
                del c2, b2
                torch.cuda.empty_cache()
                # the same metric from host buffers (copies timed) and the reference's CPU kernels beside it
                if not args.no_e2e:
                    try:
                        extras[name]["e2e"] = measure_e2e(w2, ctx, rank, world, local, 3)
                    except Exception as ex:
                        extras[name]["e2e"] = {"error": repr(ex)[:300]}
                    torch.cuda.empty_cache()
                if not args.no_e2e:
                    g = glue_timing(name, min(w2.n, 200_000_000), 3)
                    if g is not None:
                        extras[name]["glue"] = g
                if not args.no_cpu:
                    try:
                        r = run_cpu(w2, 1, 0 if w2.n > 50_000_000 else 1, host_threads())
                        extras[name]["cpu_baseline"] = {"value": r["rows_per_s"], "unit": "rows/s", "cores": r["cores"],
                                                        "kind": r["kind"], "sample": r["sample"]}
                    except Exception as ex:
                        extras[name]["cpu_baseline"] = {"value": None, "sample": "failed: %r" % (ex,)}
            except Exception as ex:  # an extra must never take the main line down
                extras[name] = {"error": repr(ex)[:300]}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        try:
            r = run_cpu(wl, 2, 1, host_threads())
            cpu = {"value": r["rows_per_s"], "unit": "rows/s", "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]}
        except Exception as ex:
            cpu = {"value": None, "unit": "rows/s", "cores": 0, "kind": "port", "sample": "failed: %r" % (ex,)}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "rows/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": wl.dtype,
            "data": "synthetic",
            "config": workload_config(wl, world),
            "input_gbs": value * wl.input_bytes() / wl.n / 1e9,
            "roofline": roof, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
            "kernel_ms_per_step": {k: v[0] / kt_steps for k, v in kt.items() if v[1]},
            "plan": getattr(wl, "plan", None),
            "kernel_timing": "per-launch CUDA events of a second pass of %d steps (the hook is off during the timed steps)" % kt_steps,
        }
        if world == 1 and not args.no_e2e:
            g = glue_timing(wl.name, min(wl.n, 200_000_000))
            if g is not None:
                line["glue"] = g
        if self_check is not None:
            line["self_check"] = self_check
        if extras:
            line["workloads"] = extras
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="sort_i64")
    ap.add_argument("--rows", type=int, default=None, help="rows per GPU (default: the config's size)")
    ap.add_argument("--no-extras", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    claim_stdout()
    if args.warmup < 3 and args.impl == "b200":
        log("bench: warmup raised to 3 (timing rules)")
        args.warmup = 3
    if args.impl == "reference":
        return reference_arm(args)
    return gpu_arm(args)


if __name__ == "__main__":
    sys.exit(main())
