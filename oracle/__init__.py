"""CPU oracle — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package.  Two checkers share one calling convention (host `Column`s in, numpy out):

  kind="port"       oracle/oracle.cc — a sequential restatement of the reference algorithm
                    (built with g++ into oracle/_build/liboracle.so)
  kind="reference"  oracle/ref_driver.cc over the UNMODIFIED reference compiled from
                    /root/reference by oracle/build_ref.sh into oracle/_ref/ (travels to the GPU
                    box as a prebuilt .so; /root/reference itself is never read at run time)
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys
from typing import List, Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)

from arrow1_b200.columns import (BINARY, BITS, BOOL, HASH_MIN_MAX, HASH_SUM, NP_OF, AggSpec, Column, agg_out_types,  # noqa: E402
                                 b200_agg_spec, b200_col, b200_out_col, col_array, unpack_bits)

PORT_SO = os.path.join(_HERE, "_build", "liboracle.so")
REF_SO = os.path.join(_HERE, "_ref", "libref_driver.so")
REF_LIB = os.path.join(_HERE, "_ref", "libarrow_ref.a")
REFERENCE_SRC = "/root/reference/cpp/src"

_libs = {}


def _run(cmd: List[str]) -> None:
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("oracle build failed: %s\n%s" % (" ".join(cmd), r.stdout[-4000:]))


def build(force: bool = False) -> None:
    """Compiles the C++ restatement and, when /root/reference is present, the reference driver."""
    src = os.path.join(_HERE, "oracle.cc")
    hdr = os.path.getmtime(os.path.join(_ROOT, "include", "arrow1_b200.h"))   # b200_col lives there
    os.makedirs(os.path.dirname(PORT_SO), exist_ok=True)
    if force or not os.path.exists(PORT_SO) or os.path.getmtime(PORT_SO) < max(os.path.getmtime(src), hdr):
        _run(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-I", os.path.join(_ROOT, "include"), src, "-o", PORT_SO])
    if os.path.isdir(REFERENCE_SRC):
        drv = os.path.join(_HERE, "ref_driver.cc")
        if force or not os.path.exists(REF_LIB):
            _run(["bash", os.path.join(_HERE, "build_ref.sh")])
        if force or not os.path.exists(REF_SO) or os.path.getmtime(REF_SO) < max(os.path.getmtime(drv), os.path.getmtime(REF_LIB), hdr):
            _run(["g++", "-std=c++11", "-O2", "-fPIC", "-shared", "-w", "-include", "limits", "-include", "cstdint",
                  "-DARROW_STATIC", "-I", os.path.join(_HERE, "_ref", "include"), "-I", REFERENCE_SRC,
                  "-I", os.path.join(_ROOT, "include"), drv, "-o", REF_SO, REF_LIB, "-lpthread", "-ldl"])


def have_reference() -> bool:
    return os.path.exists(REF_SO)


def _lib(kind: str):
    if kind in _libs:
        return _libs[kind]
    if kind == "port":
        if not os.path.exists(PORT_SO):
            build()
        lib, prefix = C.CDLL(PORT_SO), "oracle"
    elif kind == "reference":
        if not os.path.exists(REF_SO):
            raise RuntimeError("oracle/_ref/libref_driver.so not built (run oracle.build() where /root/reference exists)")
        lib, prefix = C.CDLL(REF_SO), "ref"
    else:
        raise ValueError(kind)
    _libs[kind] = (lib, prefix)
    return _libs[kind]


def _fn(kind: str, name: str):
    lib, prefix = _lib(kind)
    f = getattr(lib, "%s_%s" % (prefix, name))
    f.restype = C.c_int
    return f


class OracleError(RuntimeError):
    def __init__(self, code: int, msg: str = ""):
        super().__init__("oracle status %d %s" % (code, msg))
        self.code = code


def _check(kind: str, code: int) -> None:
    if code != 0:
        msg = ""
        if kind == "reference":
            lib, _ = _lib(kind)
            lib.ref_last_error.restype = C.c_char_p
            msg = lib.ref_last_error().decode()
        raise OracleError(code, msg)


def _out_col(type_: int, n: int, extra_cap: int = 0):
    """Allocates host output buffers for n elements; returns (struct, data, validity).  BINARY (reference only):
    `data` is the pair (int32 offsets [n + 1], value bytes [extra_cap])."""
    if type_ == BINARY:
        offs = np.zeros(n + 2, dtype=np.int32)
        blob = np.zeros(max(extra_cap, 1) + 8, dtype=np.uint8)
        validity = np.zeros((n + 7) // 8 + 8, dtype=np.uint8)
        oc = b200_out_col(validity.ctypes.data, offs.ctypes.data, type_, 0, 0, 0, blob.ctypes.data)
        return oc, (offs, blob), validity
    bits = BITS[type_]
    nbytes = (n + 7) // 8 + 8 if bits == 1 else n * (bits // 8) + 8
    data = np.zeros(nbytes, dtype=np.uint8)
    validity = np.zeros((n + 7) // 8 + 8, dtype=np.uint8)
    oc = b200_out_col(validity.ctypes.data, data.ctypes.data, type_, 0, 0, 0)
    return oc, data, validity


def _decode(type_: int, data, validity: np.ndarray, n: int):
    if type_ == BINARY:
        offs, blob = data
        valid = unpack_bits(validity, n)
        raw = blob.tobytes()
        vals = np.empty(n, dtype=object)
        for i in range(n):
            vals[i] = raw[offs[i]: offs[i + 1]] if valid[i] else b""
        return vals, valid
    if type_ == BOOL:
        vals = unpack_bits(data, n)
    else:
        dt = np.dtype(NP_OF[type_])
        vals = data[: n * dt.itemsize].view(dt).copy()
    return vals, unpack_bits(validity, n)


class FilterResult:
    def __init__(self, type_, length, null_count, has_validity, data, validity):
        self.type, self.length, self.null_count, self.has_validity = type_, length, null_count, has_validity
        self.data, self.validity = data, validity  # raw byte buffers (offset 0)

    def values(self):
        return _decode(self.type, self.data, self.validity, self.length)[0]

    def valid(self):
        if not self.has_validity:
            return np.ones(self.length, dtype=np.bool_)
        return unpack_bits(self.validity, self.length)

    def data_bytes(self) -> bytes:
        bits = BITS[self.type]
        n = (self.length + 7) // 8 if bits == 1 else self.length * (bits // 8)
        return self.data[:n].tobytes()

    def validity_bytes(self) -> Optional[bytes]:
        if not self.has_validity:
            return None
        return self.validity[: (self.length + 7) // 8].tobytes()


def filter(values: Column, mask: Column, null_selection: int = 0, kind: str = "port") -> FilterResult:
    oc, data, validity = _out_col(values.type, max(values.length, mask.length))
    has_validity = C.c_int(0)
    v, m = values.desc(), mask.desc()
    _check(kind, _fn(kind, "filter")(C.byref(v), C.byref(m), C.c_int(null_selection), C.byref(oc), C.byref(has_validity)))
    return FilterResult(values.type, oc.length, oc.null_count, bool(has_validity.value), data, validity)


def take(values: Column, indices: Column, boundscheck: bool = True, kind: str = "port") -> FilterResult:
    """compute::Take(values, indices): FilterResult-shaped (the validity bitmap is always present)."""
    oc, data, validity = _out_col(values.type, indices.length)
    v, i = values.desc(), indices.desc()
    _check(kind, _fn(kind, "take")(C.byref(v), C.byref(i), C.c_int(1 if boundscheck else 0), C.byref(oc)))
    return FilterResult(values.type, oc.length, oc.null_count, True, data, validity)


def sort_indices(values: Column, order: int = 0, index_base: int = 0, kind: str = "port") -> np.ndarray:
    out = np.zeros(max(values.length, 1), dtype=np.uint64)
    v = values.desc()
    _check(kind, _fn(kind, "sort_indices")(C.byref(v), C.c_int(order), C.c_uint64(index_base),
                                           out.ctypes.data_as(C.c_void_p)))
    return out[: values.length]


def sort_indices_multi(keys: Sequence[Column], orders: Sequence[int], kind: str = "port", as_table: bool = False,
                       n_chunks: int = 1) -> np.ndarray:
    """compute::SortIndices(RecordBatch | Table, SortOptions{sort_keys}); keys[0] is the most significant key.
    kind="reference": as_table / n_chunks choose the Table form (MultipleKeyTableSorter) over a chunked table."""
    n = keys[0].length if keys else 0
    out = np.zeros(max(n, 1), dtype=np.uint64)
    descs = (type(keys[0].desc()) * len(keys))(*[k.desc() for k in keys])
    ords = (C.c_int * len(keys))(*[int(o) for o in orders])
    if kind == "reference":
        _check(kind, _fn(kind, "sort_indices_multi")(descs, ords, C.c_int(len(keys)), C.c_int(1 if as_table else 0),
                                                     C.c_int(n_chunks), out.ctypes.data_as(C.c_void_p)))
    else:
        _check(kind, _fn(kind, "sort_indices_multi")(descs, ords, C.c_int(len(keys)), out.ctypes.data_as(C.c_void_p)))
    return out[:n]


class GroupByResult:
    """aggs: list (per requested aggregate) of list of (values, valid) numpy pairs; keys: list of
    (values, valid); ids: per-row group ids or None."""

    def __init__(self, num_groups, aggs, keys, ids):
        self.num_groups, self.aggs, self.keys, self.ids = num_groups, aggs, keys, ids


def groupby(keys: Sequence[Column], aggs: Sequence[AggSpec], args: Sequence[Column], want_ids: bool = False,
            kind: str = "port", max_groups: Optional[int] = None) -> GroupByResult:
    """`max_groups` bounds the output buffers when the caller knows the cardinality (default: n)."""
    n = keys[0].length
    cap = max(n if max_groups is None else min(n, max_groups), 1)
    specs = (b200_agg_spec * max(len(aggs), 1))()
    out_types = []
    for j, a in enumerate(aggs):
        specs[j] = b200_agg_spec(a.kind, args[j].type, a.option, 0)
        out_types.append(agg_out_types(a.kind, args[j].type))
    flat = [t for ts in out_types for t in ts]
    out_aggs = (b200_out_col * max(len(flat), 1))()
    holders = []
    for i, t in enumerate(flat):
        oc, d, v = _out_col(t, cap)
        out_aggs[i] = oc
        holders.append((t, d, v))
    out_keys = (b200_out_col * len(keys))()
    key_holders = []
    for i, k in enumerate(keys):
        oc, d, v = _out_col(k.type, cap, 0 if k.extra is None else len(k.extra))
        out_keys[i] = oc
        key_holders.append((k.type, d, v))
    ids = np.zeros(max(n, 1), dtype=np.uint32) if want_ids else None
    num_groups = C.c_uint32(0)
    _check(kind, _fn(kind, "groupby")(C.c_int(len(keys)), col_array(keys), C.c_int(len(aggs)), specs, col_array(args),
                                      out_aggs, out_keys, C.byref(num_groups),
                                      ids.ctypes.data_as(C.c_void_p) if want_ids else None))
    G = num_groups.value
    res_aggs, i = [], 0
    for ts in out_types:
        cols = []
        for _ in ts:
            t, d, v = holders[i]
            cols.append(_decode(t, d, v, G))
            i += 1
        if len(cols) == 2:  # min_max children share the first validity
            cols[1] = (cols[1][0], cols[0][1])
        res_aggs.append(cols)
    res_keys = [_decode(t, d, v, G) for (t, d, v) in key_holders]
    return GroupByResult(G, res_aggs, res_keys, None if ids is None else ids[:n])


def groupby_sum_chunked(key_chunks: Sequence[Sequence[Column]], arg_chunks: Sequence[Sequence[Column]], max_groups: int):
    """The reference's GroupBy with hash_sum aggregates over CHUNKED columns (kind="reference" only):
    key_chunks[k][c] / arg_chunks[j][c] = chunk c of key k / argument j, every chunk its own buffers at offset 0.
    This is the multi-batch form upstream computes correctly; sums accumulate in row order across the chunks
    (hash_aggregate.cc:533-545, 926-943).  Returns (num_groups, [(values, valid) per aggregate], [(values, valid) per key])."""
    n_keys, n_aggs, n_chunks = len(key_chunks), len(arg_chunks), len(key_chunks[0])
    flat_keys = [c for cs in key_chunks for c in cs]
    flat_args = [c for cs in arg_chunks for c in cs]
    for c in flat_keys + flat_args:
        assert c.offset == 0 and c.length <= (1 << 27)
    out_aggs = (b200_out_col * max(n_aggs, 1))()
    holders = []
    for j in range(n_aggs):
        t = agg_out_types(HASH_SUM, arg_chunks[j][0].type)[0]
        oc, d, v = _out_col(t, max_groups)
        out_aggs[j] = oc
        holders.append((t, d, v))
    out_keys = (b200_out_col * n_keys)()
    key_holders = []
    for k in range(n_keys):
        oc, d, v = _out_col(key_chunks[k][0].type, max_groups)
        out_keys[k] = oc
        key_holders.append((key_chunks[k][0].type, d, v))
    num_groups = C.c_uint32(0)
    _check("reference", _fn("reference", "groupby_sum_chunked")(C.c_int(n_keys), C.c_int(n_chunks), col_array(flat_keys),
                                                                  C.c_int(n_aggs), col_array(flat_args), out_aggs, out_keys,
                                                                  C.byref(num_groups)))
    G = num_groups.value
    assert G <= max_groups
    return G, [_decode(t, d, v, G) for (t, d, v) in holders], [_decode(t, d, v, G) for (t, d, v) in key_holders]


def make_groupings(ids: np.ndarray, num_groups: int):
    """Grouper::MakeGroupings of the reference (kind="reference" only): (offsets int32[num_groups + 1], indices int32[n])."""
    ids = np.ascontiguousarray(ids, dtype=np.uint32)
    offsets = np.zeros(num_groups + 1, dtype=np.int32)
    indices = np.zeros(max(len(ids), 1), dtype=np.int32)
    _check("reference", _fn("reference", "make_groupings")(ids.ctypes.data_as(C.c_void_p), C.c_int64(len(ids)), C.c_uint32(num_groups),
                                                             offsets.ctypes.data_as(C.c_void_p), indices.ctypes.data_as(C.c_void_p)))
    return offsets, indices[: len(ids)]


def value_counts(values: Column, want_counts: bool = True):
    """"unique" / "value_counts" of the reference (vector_hash.cc): ((values, valid), counts or None) in
    first-occurrence order."""
    n = values.length
    oc, d, v = _out_col(values.type, max(n, 1), 0 if values.extra is None else len(values.extra))
    counts = np.zeros(max(n, 1), dtype=np.int64)
    out_n = C.c_int64(0)
    _check("reference", _fn("reference", "value_counts")(C.byref(values.desc()), C.byref(oc),
                                                           counts.ctypes.data_as(C.c_void_p) if want_counts else None, C.byref(out_n)))
    g = out_n.value
    return _decode(values.type, d, v, g), (counts[:g].copy() if want_counts else None)


def dictionary_encode(values: Column, encode_nulls: bool = False):
    """"dictionary_encode" of the reference (vector_hash.cc:165-226): (indices int32, index validity bool,
    (dictionary values, valid))."""
    n = values.length
    idx = np.zeros(max(n, 1), dtype=np.int32)
    valid = np.zeros(max(n, 1), dtype=np.uint8)
    oc, d, v = _out_col(values.type, max(n, 1), 0 if values.extra is None else len(values.extra))
    out_n = C.c_int64(0)
    _check("reference", _fn("reference", "dictionary_encode")(C.byref(values.desc()), C.c_int(1 if encode_nulls else 0),
                                                                idx.ctypes.data_as(C.c_void_p), valid.ctypes.data_as(C.c_void_p),
                                                                C.byref(oc), C.byref(out_n)))
    return idx[:n].copy(), valid[:n].astype(bool), _decode(values.type, d, v, out_n.value)

