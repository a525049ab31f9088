"""Column descriptors shared by the C-ABI wrapper, the oracle wrappers, the tests and bench.py.

A `Column` is the Python image of `b200_col` (include/arrow1_b200.h), itself the mirror of the
reference's `PrimitiveArg` over `ArrayData` (compute/kernels/util_internal.h:35-47,
array/data.h:73-245): a values buffer, an optional LSB-first validity bitmap, an element offset,
a length and a possibly-unknown null_count.  Buffers are either numpy uint8 arrays (host) or
torch uint8 tensors (device).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional

import numpy as np

# == arrow::Type::type (type_fwd.h) == b200_type
BOOL, UINT8, INT8, UINT16, INT16, UINT32, INT32, UINT64, INT64, FLOAT, DOUBLE = 1, 2, 3, 4, 5, 6, 7, 8, 9, 11, 12
BINARY = 13   # binary / utf8 with int32 offsets: group keys only; `data` = offsets, `extra` = value bytes

NP_OF = {
    UINT8: np.uint8, INT8: np.int8, UINT16: np.uint16, INT16: np.int16, UINT32: np.uint32,
    INT32: np.int32, UINT64: np.uint64, INT64: np.int64, FLOAT: np.float32, DOUBLE: np.float64,
}
TYPE_OF = {np.dtype(v): k for k, v in NP_OF.items()}
TYPE_OF[np.dtype(np.bool_)] = BOOL
BITS = {BOOL: 1, UINT8: 8, INT8: 8, UINT16: 16, INT16: 16, UINT32: 32, INT32: 32, UINT64: 64,
        INT64: 64, FLOAT: 32, DOUBLE: 64}
NAMES = {BOOL: "bool", UINT8: "uint8", INT8: "int8", UINT16: "uint16", INT16: "int16",
         UINT32: "uint32", INT32: "int32", UINT64: "uint64", INT64: "int64", FLOAT: "float",
         DOUBLE: "double", BINARY: "binary"}

# aggregate kinds == b200_agg_kind
HASH_COUNT, HASH_SUM, HASH_MIN_MAX, HASH_MEAN = 0, 1, 2, 3
AGG_NAMES = {HASH_COUNT: "hash_count", HASH_SUM: "hash_sum", HASH_MIN_MAX: "hash_min_max", HASH_MEAN: "hash_mean"}


class b200_col(C.Structure):
    _fields_ = [("validity", C.c_void_p), ("data", C.c_void_p), ("type", C.c_int32),
                ("reserved", C.c_int32), ("length", C.c_int64), ("offset", C.c_int64),
                ("null_count", C.c_int64), ("extra", C.c_void_p)]


class b200_out_col(C.Structure):
    _fields_ = [("validity", C.c_void_p), ("data", C.c_void_p), ("type", C.c_int32),
                ("reserved", C.c_int32), ("length", C.c_int64), ("null_count", C.c_int64), ("extra", C.c_void_p)]


class b200_agg_spec(C.Structure):
    _fields_ = [("kind", C.c_int32), ("arg_type", C.c_int32), ("option", C.c_int32),
                ("reserved", C.c_int32)]


class b200_fused_arg(C.Structure):
    _fields_ = [("kind", C.c_int32), ("op1", C.c_int32), ("op2", C.c_int32), ("reserved", C.c_int32),
                ("a", b200_col), ("b", b200_col), ("scalar", C.c_double)]


class b200_predicate(C.Structure):
    _fields_ = [("column", b200_col), ("op", C.c_int32), ("reserved", C.c_int32), ("scalar", C.c_void_p)]


def pack_bits(bits: np.ndarray, pad_to: int = 8) -> np.ndarray:
    """bool array -> LSB-first bitmap (uint8), zero padded to a multiple of `pad_to` bytes."""
    packed = np.packbits(np.asarray(bits, dtype=np.bool_), bitorder="little")
    n = (len(packed) + pad_to - 1) // pad_to * pad_to
    out = np.zeros(max(n, pad_to), dtype=np.uint8)
    out[: len(packed)] = packed
    return out


def unpack_bits(buf: np.ndarray, length: int, offset: int = 0) -> np.ndarray:
    return np.unpackbits(np.asarray(buf, dtype=np.uint8), bitorder="little")[offset: offset + length].astype(np.bool_)


def _ptr(buf) -> Optional[int]:
    if buf is None:
        return None
    if isinstance(buf, np.ndarray):
        return buf.ctypes.data
    return buf.data_ptr()  # torch tensor


@dataclass
class Column:
    type: int
    data: object                # numpy uint8 / torch uint8 buffer holding offset+length elements
    validity: Optional[object]  # bitmap buffer or None
    length: int
    offset: int = 0
    null_count: int = -1
    extra: Optional[object] = None   # BINARY: the value bytes (`data` holds the int32 offsets)

    # ---- constructors ------------------------------------------------------------------
    @staticmethod
    def from_numpy(values: np.ndarray, valid: Optional[np.ndarray] = None, offset: int = 0,
                   null_count: Optional[int] = None, type: Optional[int] = None) -> "Column":
        """`values` (and `valid`) describe the PARENT buffers; the column is parent[offset:]."""
        values = np.ascontiguousarray(values)
        t = type if type is not None else TYPE_OF[values.dtype]
        n = len(values) - offset
        if t == BOOL:
            data = pack_bits(values.astype(np.bool_))
        else:
            raw = values.view(np.uint8).reshape(-1)
            data = np.zeros((len(raw) + 63) // 64 * 64 + 64, dtype=np.uint8)
            data[: len(raw)] = raw
        vbuf = None
        nc = 0
        if valid is not None:
            valid = np.asarray(valid, dtype=np.bool_)
            assert len(valid) == len(values)
            vbuf = pack_bits(valid)
            nc = int((~valid[offset:]).sum())
        if null_count is not None:
            nc = null_count
        return Column(t, data, vbuf, n, offset, nc)

    @staticmethod
    def from_strings(items, offset: int = 0) -> "Column":
        """A BINARY column (int32 offsets) from bytes / str / None entries; the column is parent[offset:]."""
        raw = [None if x is None else (x.encode() if isinstance(x, str) else bytes(x)) for x in items]
        lens = np.array([0 if x is None else len(x) for x in raw], dtype=np.int64)
        offs = np.zeros(len(raw) + 1, dtype=np.int32)
        np.cumsum(lens, out=offs[1:])
        blob = b"".join(x for x in raw if x is not None)
        data = np.zeros((offs.nbytes + 63) // 64 * 64 + 64, dtype=np.uint8)
        data[: offs.nbytes] = offs.view(np.uint8)
        extra = np.zeros((len(blob) + 63) // 64 * 64 + 64, dtype=np.uint8)
        extra[: len(blob)] = np.frombuffer(blob, dtype=np.uint8)
        valid = np.array([x is not None for x in raw], dtype=np.bool_)
        vbuf = None if valid.all() else pack_bits(valid)
        return Column(BINARY, data, vbuf, len(raw) - offset, offset, int((~valid[offset:]).sum()), extra)

    def strings(self):
        """BINARY column -> list of bytes / None."""
        h = self.to_host()
        offs = h.data[: (h.offset + h.length + 1) * 4].view(np.int32)[h.offset:]
        valid = h.valid()
        blob = h.extra.tobytes() if h.extra is not None else b""
        return [blob[offs[i]: offs[i + 1]] if valid[i] else None for i in range(h.length)]

    # ---- views -------------------------------------------------------------------------
    def desc(self) -> b200_col:
        return b200_col(_ptr(self.validity), _ptr(self.data), self.type, 0, self.length, self.offset,
                        self.null_count, _ptr(self.extra))

    @property
    def on_device(self) -> bool:
        return not isinstance(self.data, np.ndarray)

    def to_device(self, device="cuda") -> "Column":
        import torch
        d = torch.from_numpy(self.data).to(device)
        v = None if self.validity is None else torch.from_numpy(self.validity).to(device)
        x = None if self.extra is None else torch.from_numpy(self.extra).to(device)
        return Column(self.type, d, v, self.length, self.offset, self.null_count, x)

    def to_host(self) -> "Column":
        if not self.on_device:
            return self
        d = self.data.cpu().numpy()
        v = None if self.validity is None else self.validity.cpu().numpy()
        x = None if self.extra is None else self.extra.cpu().numpy()
        return Column(self.type, d, v, self.length, self.offset, self.null_count, x)

    def values(self) -> np.ndarray:
        """Logical values (host) as a numpy array of the column type (bool for BOOL)."""
        h = self.to_host()
        if h.type == BOOL:
            return unpack_bits(h.data, h.length, h.offset)
        dt = np.dtype(NP_OF[h.type])
        a = h.data[: (h.offset + h.length) * dt.itemsize].view(dt)
        return a[h.offset: h.offset + h.length]

    def valid(self) -> np.ndarray:
        h = self.to_host()
        if h.validity is None:
            return np.ones(h.length, dtype=np.bool_)
        return unpack_bits(h.validity, h.length, h.offset)

    def nbytes(self) -> int:
        n = self.data.nbytes if isinstance(self.data, np.ndarray) else self.data.numel()
        if self.validity is not None:
            n += self.validity.nbytes if isinstance(self.validity, np.ndarray) else self.validity.numel()
        return n


def col_array(cols) -> "C.Array":
    arr = (b200_col * max(len(cols), 1))()
    for i, c in enumerate(cols):
        arr[i] = c.desc()
    return arr


@dataclass
class AggSpec:
    kind: int
    option: int = 0

    def name(self) -> str:
        return AGG_NAMES[self.kind]


def agg_out_types(kind: int, arg_type: int):
    """Output column types of one aggregate, in b200_groupby_finalize order."""
    if kind == HASH_COUNT:
        return [INT64]
    if kind == HASH_MEAN:
        return [DOUBLE]
    if kind == HASH_SUM:
        # FindAccumulatorType (kernels/aggregate_internal.h:30-50)
        if arg_type in (FLOAT, DOUBLE):
            return [DOUBLE]
        if arg_type in (INT8, INT16, INT32, INT64):
            return [INT64]
        return [UINT64]
    if kind == HASH_MIN_MAX:
        return [arg_type, arg_type]
    raise ValueError(kind)
