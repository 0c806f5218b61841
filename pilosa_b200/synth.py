"""Seeded synthetic index generator (SURVEY.md §8d workloads C1..C5): ctypes binding of
libpb200_synth.so. Every generator returns the bytes of a Pilosa fragment file, so the same
bytes feed the engine (pb200_frag_load) and, in tests / the CPU baseline, the oracle.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Sequence

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libpb200_synth.so")
DEFAULT_SEED = 20190416

_lib = None


def _load() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            raise ImportError(f"{_SO} is missing: run `python __graft_entry__.py` to build it")
        L = C.CDLL(_SO)
        vp = C.c_void_p
        L.pbs_free.argtypes = [vp]
        L.pbs_data.restype, L.pbs_data.argtypes = vp, [vp]
        L.pbs_size.restype, L.pbs_size.argtypes = C.c_size_t, [vp]
        L.pbs_gen_bernoulli.restype = vp
        L.pbs_gen_bernoulli.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_double]
        L.pbs_gen_bernoulli_rows.restype = vp
        L.pbs_gen_bernoulli_rows.argtypes = [C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64), C.c_uint64, C.c_double]
        L.pbs_gen_mixed.restype = vp
        L.pbs_gen_mixed.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64]
        L.pbs_gen_bsi.restype = vp
        L.pbs_gen_bsi.argtypes = [C.c_uint64, C.c_uint64, C.c_uint32, C.c_int64, C.c_int64, C.c_double]
        L.pbs_gen_zipf.restype = vp
        L.pbs_gen_zipf.argtypes = [C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32, C.c_double, C.c_double]
        L.pbs_from_positions.restype = vp
        L.pbs_from_positions.argtypes = [C.POINTER(C.c_uint64), C.c_size_t]
        _lib = L
    return _lib


class FragBytes:
    """Owns a generated fragment image; ``ptr``/``size`` can be handed to the C ABI without
    a copy, ``bytes()`` materialises a Python bytes object."""

    def __init__(self, handle):
        self._h = handle
        L = _load()
        self.ptr = L.pbs_data(handle)
        self.size = L.pbs_size(handle)

    def bytes(self) -> bytes:
        return C.string_at(self.ptr, self.size)

    def __len__(self) -> int:
        return self.size

    def __del__(self):
        if getattr(self, "_h", None):
            _load().pbs_free(self._h)
            self._h = None


def bernoulli(shard: int, row_lo: int, n_rows: int, p: float, seed: int = DEFAULT_SEED) -> FragBytes:
    return FragBytes(_load().pbs_gen_bernoulli(seed, shard, row_lo, n_rows, p))


def bernoulli_rows(shard: int, row_ids: Sequence[int], p: float, seed: int = DEFAULT_SEED) -> FragBytes:
    """Same per-(shard,row,container) streams as ``bernoulli`` for an explicit ascending row list."""
    ids = sorted(row_ids)
    arr = (C.c_uint64 * max(1, len(ids)))(*ids)
    return FragBytes(_load().pbs_gen_bernoulli_rows(seed, shard, arr, len(ids), p))


def mixed(shard: int, row_lo: int, n_rows: int, seed: int = DEFAULT_SEED) -> FragBytes:
    return FragBytes(_load().pbs_gen_mixed(seed, shard, row_lo, n_rows))


def bsi(shard: int, bit_depth: int = 32, vmin: int = -(1 << 31), vmax: int = (1 << 31) - 1, density: float = 1.0,
        seed: int = DEFAULT_SEED) -> FragBytes:
    return FragBytes(_load().pbs_gen_bsi(seed, shard, bit_depth, vmin, vmax, density))


def zipf(shard: int, n_rows: int = 10000, k: int = 8, s: float = 1.6, v: float = 50.0, seed: int = DEFAULT_SEED) -> FragBytes:
    return FragBytes(_load().pbs_gen_zipf(seed, shard, n_rows, k, s, v))


def from_positions(positions: Sequence[int]) -> FragBytes:
    arr = (C.c_uint64 * max(1, len(positions)))(*positions)
    return FragBytes(_load().pbs_from_positions(arr, len(positions)))
