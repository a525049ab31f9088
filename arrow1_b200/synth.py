"""Counter-based synthetic columns of the benchmark configs (DESIGN.md "Inputs").

    r(s, i) = splitmix64(s * 0x9E3779B97F4A7C15 + i)          u(s, i) = (r >> 11) * 2^-53

Every row depends only on (seed, global row index), so each GPU rank and the CPU baseline
generate identical bytes for the same rows with no transfer.  Two back ends produce the same
bytes: numpy (host) and torch (device); torch is used as an array calculator only.
"""
from __future__ import annotations

import numpy as np

from . import columns as cols
from .columns import Column

GOLDEN = 0x9E3779B97F4A7C15
_M1, _M2 = 0xBF58476D1CE4E5B9, 0x94D049BB133111EB
_MASK = (1 << 64) - 1


def _s64(x: int) -> int:
    x &= _MASK
    return x - (1 << 64) if x >= (1 << 63) else x


# ---- numpy -------------------------------------------------------------------------------------
def r_np(seed: int, start: int, n: int) -> np.ndarray:
    with np.errstate(over="ignore"):
        x = np.arange(start, start + n, dtype=np.uint64) + np.uint64((seed * GOLDEN) & _MASK)
        x = x + np.uint64(GOLDEN)
        z = (x ^ (x >> np.uint64(30))) * np.uint64(_M1)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(_M2)
        return z ^ (z >> np.uint64(31))


def u_np(seed: int, start: int, n: int) -> np.ndarray:
    return (r_np(seed, start, n) >> np.uint64(11)).astype(np.float64) * 2.0 ** -53


# ---- torch -------------------------------------------------------------------------------------
def _lsr(x, k: int):
    return (x >> k) & ((1 << (64 - k)) - 1)


def r_t(seed: int, start: int, n: int, device):
    import torch
    x = torch.arange(start, start + n, dtype=torch.int64, device=device) + _s64(seed * GOLDEN)
    x = x + _s64(GOLDEN)
    z = (x ^ _lsr(x, 30)) * _s64(_M1)
    z = (z ^ _lsr(z, 27)) * _s64(_M2)
    return z ^ _lsr(z, 31)


def r_at(seed: int, idx):
    """r(seed, i) for an int64 tensor of row numbers (the multi-GPU self-check regenerates the keys of the rows a rank
    received from the other ranks)."""
    x = idx + _s64(seed * GOLDEN)
    x = x + _s64(GOLDEN)
    z = (x ^ _lsr(x, 30)) * _s64(_M1)
    z = (z ^ _lsr(z, 27)) * _s64(_M2)
    return z ^ _lsr(z, 31)


def u_t(seed: int, start: int, n: int, device):
    import torch
    return _lsr(r_t(seed, start, n, device), 11).to(torch.float64) * 2.0 ** -53


def _pack_t(bits):
    """bool tensor -> LSB-first bitmap (uint8 tensor), padded to 64 bytes."""
    import torch
    n = bits.numel()
    nb = (n + 7) // 8
    padded = torch.zeros(((nb + 63) // 64 * 64 + 64) * 8, dtype=torch.uint8, device=bits.device)
    padded[:n] = bits.to(torch.uint8)
    w = torch.tensor([1, 2, 4, 8, 16, 32, 64, 128], dtype=torch.uint8, device=bits.device)
    return (padded.view(-1, 8) * w).sum(dim=1, dtype=torch.uint8)


_CHUNK = 1 << 26


class Gen:
    """`device=None` -> numpy buffers; otherwise torch buffers on that device."""

    def __init__(self, device=None):
        self.device = device

    def _chunks(self, start, n, fn_np, fn_t, dtype_np, dtype_t):
        if self.device is None:
            out = np.empty(n, dtype=dtype_np)
            for o in range(0, n, _CHUNK):
                m = min(_CHUNK, n - o)
                out[o:o + m] = fn_np(start + o, m)
            return out
        import torch
        out = torch.empty(n, dtype=dtype_t, device=self.device)
        for o in range(0, n, _CHUNK):
            m = min(_CHUNK, n - o)
            out[o:o + m] = fn_t(start + o, m)
        return out

    def _bitmap(self, start, n, fn_np, fn_t):
        """fn_* return bool arrays; result is a padded bitmap buffer (chunks are multiples of 8)."""
        if self.device is None:
            bits = np.empty(n, dtype=np.bool_)
            for o in range(0, n, _CHUNK):
                m = min(_CHUNK, n - o)
                bits[o:o + m] = fn_np(start + o, m)
            return cols.pack_bits(bits, 64), int(n - bits.sum())
        import torch
        nb = (n + 7) // 8
        out = torch.zeros((nb + 63) // 64 * 64 + 64, dtype=torch.uint8, device=self.device)
        zeros = 0
        for o in range(0, n, _CHUNK):
            m = min(_CHUNK, n - o)
            b = fn_t(start + o, m)
            zeros += int(m - b.sum().item())
            p = _pack_t(b)
            k = (m + 7) // 8
            out[o // 8: o // 8 + k] = p[:k]
        return out, zeros

    def _wrap(self, t, data, validity, n, null_count):
        if self.device is None:
            raw = data.view(np.uint8).reshape(-1) if data.dtype != np.uint8 else data
            buf = np.zeros((len(raw) + 63) // 64 * 64 + 64, dtype=np.uint8)
            buf[: len(raw)] = raw
            return Column(t, buf, validity, n, 0, null_count)
        import torch
        raw = data.view(torch.uint8).reshape(-1) if data.dtype != torch.uint8 else data
        return Column(t, raw, validity, n, 0, null_count)

    # ---- validity: (r >> 33) % 100 != 0  (~1 % nulls)
    def validity(self, seed, start, n):
        return self._bitmap(start, n,
                            lambda s, m: ((r_np(seed, s, m) >> np.uint64(33)) % np.uint64(100)) != 0,
                            lambda s, m: (_lsr(r_t(seed, s, m, self.device), 33) % 100) != 0)

    # ---- C1: Filter, int64 values (seed 1), 1 % nulls (seed 2), 50 % mask (seed 3, no nulls)
    def filter_c1(self, n, start=0):
        import torch
        vals = self._chunks(start, n, lambda s, m: r_np(1, s, m).view(np.int64),
                            lambda s, m: r_t(1, s, m, self.device), np.int64, None if self.device is None else torch.int64)
        vbits, nulls = self.validity(2, start, n)
        mbits, _ = self._bitmap(start, n, lambda s, m: (r_np(3, s, m) & np.uint64(1)) != 0,
                                lambda s, m: (r_t(3, s, m, self.device) & 1) != 0)
        return self._wrap(cols.INT64, vals, vbits, n, nulls), Column(cols.BOOL, mbits, None, n, 0, 0)

    # ---- C2: SortIndices, int64 full range (seed 4) / float64 in [0, 1e14) (seed 5), 1 % nulls (seed 6)
    def sort_c2(self, n, dtype="int64", start=0):
        import torch
        if dtype == "int64":
            vals = self._chunks(start, n, lambda s, m: r_np(4, s, m).view(np.int64),
                                lambda s, m: r_t(4, s, m, self.device), np.int64, None if self.device is None else torch.int64)
            t = cols.INT64
        else:
            vals = self._chunks(start, n, lambda s, m: u_np(5, s, m) * 1e14,
                                lambda s, m: u_t(5, s, m, self.device) * 1e14, np.float64,
                                None if self.device is None else torch.float64)
            t = cols.DOUBLE
        vbits, nulls = self.validity(6, start, n)
        return self._wrap(t, vals, vbits, n, nulls)

    # ---- C3: (int32 key in [0, n_keys) seed 7, float64 value in [0, 1e14) seed 8, 1 % value nulls seed 11)
    def groupby_c3(self, n, n_keys=1000, start=0):
        import torch
        keys = self._chunks(start, n, lambda s, m: ((r_np(7, s, m) >> np.uint64(20)) % np.uint64(n_keys)).astype(np.int32),
                            lambda s, m: (_lsr(r_t(7, s, m, self.device), 20) % n_keys).to(torch.int32), np.int32,
                            None if self.device is None else torch.int32)
        vals = self._chunks(start, n, lambda s, m: u_np(8, s, m) * 1e14, lambda s, m: u_t(8, s, m, self.device) * 1e14,
                            np.float64, None if self.device is None else torch.float64)
        vbits, nulls = self.validity(11, start, n)
        return self._wrap(cols.INT32, keys, None, n, 0), self._wrap(cols.DOUBLE, vals, vbits, n, nulls)

    # ---- C4: (int64 key = mix(r % n_keys) seed 9, float64 value seed 10), no nulls
    def groupby_c4(self, n, n_keys=100_000_000, start=0):
        import torch

        def knp(s, m):
            k = (r_np(9, s, m) >> np.uint64(20)) % np.uint64(n_keys)
            with np.errstate(over="ignore"):
                z = (k + np.uint64(1)) * np.uint64(_M1)
                z = (z ^ (z >> np.uint64(29))) * np.uint64(_M2)
                return (z ^ (z >> np.uint64(32))).view(np.int64)

        def kt(s, m):
            k = _lsr(r_t(9, s, m, self.device), 20) % n_keys
            z = (k + 1) * _s64(_M1)
            z = (z ^ _lsr(z, 29)) * _s64(_M2)
            return z ^ _lsr(z, 32)

        keys = self._chunks(start, n, knp, kt, np.int64, None if self.device is None else torch.int64)
        vals = self._chunks(start, n, lambda s, m: u_np(10, s, m) * 1e14, lambda s, m: u_t(10, s, m, self.device) * 1e14,
                            np.float64, None if self.device is None else torch.float64)
        return self._wrap(cols.INT64, keys, None, n, 0), self._wrap(cols.DOUBLE, vals, None, n, 0)

    # ---- C5: Q1-shaped lineitem: shipdate date32 in [8036, 10561] seed 12; returnflag /
    # linestatus as uint8 derived from it (seed 13 picks 'A' or 'R'); quantity seed 14, extendedprice seed 15,
    # discount seed 16; all float64, no nulls
    def q1_lineitem(self, n, start=0):
        import torch
        dev = self.device
        i32 = None if dev is None else torch.int32
        u8 = None if dev is None else torch.uint8
        f64 = None if dev is None else torch.float64

        def ship_np(s, m):
            return (8036 + ((r_np(12, s, m) >> np.uint64(20)) % np.uint64(2526))).astype(np.int32)

        def ship_t(s, m):
            return (8036 + _lsr(r_t(12, s, m, dev), 20) % 2526).to(torch.int32)

        ship = self._chunks(start, n, ship_np, ship_t, np.int32, i32)

        def rf_np(s, m):
            sd = ship_np(s, m)
            ar = np.where((r_np(13, s, m) & np.uint64(1)) != 0, 65, 82)
            return np.where(sd > 9268, 78, ar).astype(np.uint8)

        def rf_t(s, m):
            sd = ship_t(s, m)
            ar = torch.where((r_t(13, s, m, dev) & 1) != 0, 65, 82)
            return torch.where(sd > 9268, 78, ar).to(torch.uint8)

        rflag = self._chunks(start, n, rf_np, rf_t, np.uint8, u8)
        lstat = self._chunks(start, n, lambda s, m: np.where(ship_np(s, m) <= 9298, 70, 79).astype(np.uint8),
                             lambda s, m: torch.where(ship_t(s, m) <= 9298, 70, 79).to(torch.uint8), np.uint8, u8)
        qty = self._chunks(start, n, lambda s, m: (1 + (r_np(14, s, m) >> np.uint64(20)) % np.uint64(50)).astype(np.float64),
                           lambda s, m: (1 + _lsr(r_t(14, s, m, dev), 20) % 50).to(torch.float64), np.float64, f64)
        price = self._chunks(start, n, lambda s, m: u_np(15, s, m) * 104100.0 + 900.0,
                             lambda s, m: u_t(15, s, m, dev) * 104100.0 + 900.0, np.float64, f64)
        disc = self._chunks(start, n, lambda s, m: ((r_np(16, s, m) >> np.uint64(20)) % np.uint64(11)).astype(np.float64) / 100.0,
                            lambda s, m: (_lsr(r_t(16, s, m, dev), 20) % 11).to(torch.float64) / 100.0, np.float64, f64)
        return (self._wrap(cols.INT32, ship, None, n, 0), self._wrap(cols.UINT8, rflag, None, n, 0),
                self._wrap(cols.UINT8, lstat, None, n, 0), self._wrap(cols.DOUBLE, qty, None, n, 0),
                self._wrap(cols.DOUBLE, price, None, n, 0), self._wrap(cols.DOUBLE, disc, None, n, 0))
