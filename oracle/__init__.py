"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes binding of oracle/_build/liboracle.so, the CPU restatement of the reference's
roaring / fragment / executor algorithms (see oracle/roaring.hpp for the header).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module; the product package ``pilosa_b200`` never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import List, Optional, Sequence, Tuple

from pilosa_b200.abi import Node, Pair, u64_array  # struct layouts only (interface)

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle.so")

NIL, ARRAY, BITMAP, RUN = 0, 1, 2, 3
OP_INTERSECT, OP_UNION, OP_DIFFERENCE, OP_XOR, OP_FLIP, OP_OPTIMIZE, OP_CONVERT, OP_SHIFT = range(8)


def build(force: bool = False) -> str:
    srcs = [os.path.join(_HERE, f) for f in ("capi.cpp", "roaring.hpp", "pilosa.hpp")]
    srcs.append(os.path.join(_HERE, "..", "include", "pilosa_b200.h"))
    stale = force or not os.path.exists(_SO) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs)
    if stale:
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    try:
        path = build()
        L = C.CDLL(path)
    except (OSError, subprocess.CalledProcessError):
        path = build(force=True)
        L = C.CDLL(path)
    u16p, u64p, i64p, ip = C.POINTER(C.c_uint16), C.POINTER(C.c_uint64), C.POINTER(C.c_int64), C.POINTER(C.c_int)
    vp, sz = C.c_void_p, C.c_size_t
    np_ = C.POINTER(Node)
    sig = {
        "orc_last_error": (C.c_char_p, []),
        "orc_container_op": (C.c_int, [C.c_int, C.c_int, u16p, C.c_int, C.c_int, u16p, C.c_int, ip, ip, u16p, ip, ip]),
        "orc_container_icount": (C.c_int, [C.c_int, u16p, C.c_int, C.c_int, u16p, C.c_int, ip]),
        "orc_container_stat": (C.c_int, [C.c_int, C.c_int, u16p, C.c_int, C.c_int, C.c_int, ip]),
        "orc_bitmap_new": (vp, [u64p, sz]),
        "orc_bitmap_free": (None, [vp]),
        "orc_bitmap_add": (C.c_int, [vp, C.c_uint64]),
        "orc_bitmap_remove": (C.c_int, [vp, C.c_uint64]),
        "orc_bitmap_optimize": (None, [vp]),
        "orc_bitmap_count": (C.c_uint64, [vp]),
        "orc_bitmap_op": (vp, [C.c_int, vp, vp]),
        "orc_bitmap_union_many": (vp, [vp, C.POINTER(vp), C.c_int]),
        "orc_bitmap_intersection_count": (C.c_uint64, [vp, vp]),
        "orc_bitmap_slice": (sz, [vp, u64p, sz]),
        "orc_bitmap_containers": (sz, [vp, u64p, ip, ip, sz]),
        "orc_bitmap_serialize": (sz, [vp, C.c_int, C.c_char_p, sz]),
        "orc_bitmap_deserialize": (vp, [C.c_char_p, sz]),
        "orc_encode_op": (sz, [C.c_int, u64p, sz, C.c_char_p, sz]),
        "orc_row_new": (vp, [u64p, sz]),
        "orc_row_free": (None, [vp]),
        "orc_row_op": (vp, [C.c_int, vp, vp]),
        "orc_row_count": (C.c_uint64, [vp]),
        "orc_row_intersection_count": (C.c_uint64, [vp, vp]),
        "orc_row_columns": (sz, [vp, u64p, sz]),
        "orc_new": (vp, [C.c_int]),
        "orc_free": (None, [vp]),
        "orc_set_workers": (None, [vp, C.c_int]),
        "orc_family_create": (C.c_uint32, [vp]),
        "orc_frag_load": (C.c_int, [vp, C.c_uint32, C.c_uint64, C.c_char_p, sz]),
        "orc_frag_set_bit": (C.c_int, [vp, C.c_uint32, C.c_uint64, C.c_uint64, C.c_uint64]),
        "orc_frag_clear_bit": (C.c_int, [vp, C.c_uint32, C.c_uint64, C.c_uint64, C.c_uint64]),
        "orc_frag_set_value": (C.c_int, [vp, C.c_uint32, C.c_uint64, C.c_uint64, C.c_uint32, C.c_int64]),
        "orc_frag_clear_value": (C.c_int, [vp, C.c_uint32, C.c_uint64, C.c_uint64, C.c_uint32, C.c_int64]),
        "orc_frag_value": (C.c_int, [vp, C.c_uint32, C.c_uint64, C.c_uint64, C.c_uint32, i64p]),
        "orc_frag_cache_config": (None, [vp, C.c_uint32, C.c_uint64, C.c_int, C.c_uint32]),
        "orc_frag_recalculate_cache": (None, [vp, C.c_uint32, C.c_uint64]),
        "orc_frag_cache_add": (None, [vp, C.c_uint32, C.c_uint64, C.c_uint64, C.c_uint64]),
        "orc_frag_cache_len": (sz, [vp, C.c_uint32, C.c_uint64]),
        "orc_frag_rebuild_cache": (None, [vp, C.c_uint32, C.c_uint64]),
        "orc_frag_rankings": (sz, [vp, C.c_uint32, C.c_uint64, u64p, u64p, sz]),
        "orc_frag_serialize": (sz, [vp, C.c_uint32, C.c_uint64, C.c_char_p, sz]),
        "orc_count": (C.c_int, [vp, np_, C.c_uint32, u64p, C.c_uint32, C.c_int, u64p]),
        "orc_count_shards": (C.c_int, [vp, np_, C.c_uint32, u64p, C.c_uint32, C.c_int, u64p]),
        "orc_count_batch": (C.c_int, [vp, np_, C.POINTER(C.c_uint32), C.c_uint32, u64p, C.c_uint32, C.c_int, u64p]),
        "orc_eval_columns": (C.c_int, [vp, np_, C.c_uint32, C.c_uint64, u64p, sz, C.POINTER(sz)]),
        "orc_eval_roaring": (C.c_int, [vp, np_, C.c_uint32, C.c_uint64, C.c_char_p, sz, C.POINTER(sz)]),
        "orc_eval_proto_columns": (C.c_int, [vp, np_, C.c_uint32, u64p, C.c_uint32, C.c_char_p, sz, C.POINTER(sz)]),
        "orc_frag_drop": (C.c_int, [vp, C.c_uint32, C.c_uint64]),
        "orc_frag_import_roaring": (C.c_int, [vp, C.c_uint32, C.c_uint64, C.c_char_p, sz, C.c_int, C.POINTER(C.c_int64)]),
        "orc_frag_import_positions": (C.c_int, [vp, C.c_uint32, C.c_uint64, u64p, sz, u64p, sz, C.POINTER(C.c_int64)]),
        "orc_rows": (C.c_int, [vp, C.c_uint32, C.c_uint64, C.c_uint32, C.c_uint64, C.c_uint64, u64p, C.c_uint32, u64p, sz, C.POINTER(sz)]),
        "orc_groupby": (C.c_int, [vp, C.c_void_p, C.c_uint32, np_, C.c_uint32, C.c_uint64, u64p, C.c_uint32, C.c_void_p, sz, C.POINTER(sz)]),
        "orc_topn": (C.c_int, [vp, C.c_uint32, np_, C.c_uint32, C.c_uint32, u64p, sz, C.c_uint64, C.c_uint64, C.c_uint32, u64p,
                               C.c_uint32, C.POINTER(Pair), sz, C.POINTER(sz)]),
        "orc_topn_shard": (C.c_int, [vp, C.c_uint32, np_, C.c_uint32, C.c_uint32, u64p, sz, C.c_uint64, C.c_uint64,
                                     C.c_uint64, C.POINTER(Pair), sz, C.POINTER(sz)]),
        "orc_bsi_sum": (C.c_int, [vp, C.c_uint32, C.c_uint32, np_, C.c_uint32, C.c_uint32, u64p, C.c_uint32, i64p, u64p]),
        "orc_bsi_minmax": (C.c_int, [vp, C.c_uint32, C.c_uint32, np_, C.c_uint32, C.c_int, u64p, C.c_uint32, i64p, u64p]),
        "orc_bsi_minmax_shards": (C.c_int, [vp, C.c_uint32, C.c_uint32, np_, C.c_uint32, C.c_int, u64p, C.c_uint32, i64p, u64p]),
        "orc_base_value": (C.c_int64, [C.c_int64, C.c_uint32, C.c_uint32, C.c_int64, ip]),
        "orc_base_value_between": (C.c_int, [C.c_int64, C.c_uint32, C.c_int64, C.c_int64, i64p, i64p]),
        "orc_bit_depth": (C.c_uint32, [C.c_uint64]),
        "orc_bit_depth_i64": (C.c_uint32, [C.c_int64]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype, fn.argtypes = res, args
    _lib = L
    return L


class OracleError(RuntimeError):
    pass


def _check(rc: int) -> None:
    if rc != 0:
        raise OracleError(lib().orc_last_error().decode())


def _u16(vals: Sequence[int]):
    return (C.c_uint16 * max(1, len(vals)))(*vals)


# ---------------------------------------------------------------------------------------
# Container-level helpers (golden vectors of roaring_internal_test.go)
# ---------------------------------------------------------------------------------------
class Cont:
    """A container literal: type + payload (array/bitmap: values; run: (start,last) pairs)."""

    def __init__(self, typ: int, payload: Sequence):
        self.typ = typ
        if typ == RUN:
            flat: List[int] = []
            for s, l in payload:
                flat += [s, l]
            self.flat, self.n = flat, len(payload)
        else:
            self.flat, self.n = list(payload), len(payload)
        self.c = (C.c_uint16 * max(1, len(self.flat)))(*self.flat)  # cached ctypes view

    @staticmethod
    def array(vals):
        return Cont(ARRAY, vals)

    @staticmethod
    def bitmap(vals):
        return Cont(BITMAP, vals)

    @staticmethod
    def bitmap_words(words: Sequence[int]):
        vals = [i * 64 + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1]
        return Cont(BITMAP, vals)

    @staticmethod
    def run(pairs):
        return Cont(RUN, pairs)

    @staticmethod
    def nil():
        return Cont(NIL, [])


def container_op(op: int, a: Cont, b: Optional[Cont] = None, to_type: int = 0) -> Tuple[int, List[int], int, bool]:
    """Returns (result type, sorted values, representation length, carry)."""
    b = b or Cont.nil()
    out_type, out_n, rep, carry = C.c_int(), C.c_int(), C.c_int(), C.c_int()
    out = (C.c_uint16 * 65536)()
    tb = to_type if op == OP_CONVERT else b.typ
    _check(lib().orc_container_op(op, a.typ, a.c, a.n, tb, b.c, b.n, C.byref(out_type),
                                  C.byref(out_n), out, C.byref(rep), C.byref(carry)))
    return out_type.value, list(out[: out_n.value]), rep.value, bool(carry.value)


def container_icount(a: Cont, b: Cont) -> int:
    out = C.c_int()
    _check(lib().orc_container_icount(a.typ, a.c, a.n, b.typ, b.c, b.n, C.byref(out)))
    return out.value


def container_stat(what: int, a: Cont, start: int = 0, end: int = 0) -> int:
    out = C.c_int()
    _check(lib().orc_container_stat(what, a.typ, a.c, a.n, start, end, C.byref(out)))
    return out.value


# ---------------------------------------------------------------------------------------
# Bitmap / Row level
# ---------------------------------------------------------------------------------------
class Bitmap:
    def __init__(self, *vals: int, _h=None):
        self._h = _h if _h is not None else lib().orc_bitmap_new(u64_array(vals), len(vals))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_bitmap_free(self._h)
            self._h = None

    def add(self, v: int) -> bool:
        return bool(lib().orc_bitmap_add(self._h, v))

    def remove(self, v: int) -> bool:
        return bool(lib().orc_bitmap_remove(self._h, v))

    def optimize(self) -> None:
        lib().orc_bitmap_optimize(self._h)

    def count(self) -> int:
        return lib().orc_bitmap_count(self._h)

    def _op(self, op: int, other: Optional["Bitmap"]) -> "Bitmap":
        return Bitmap(_h=lib().orc_bitmap_op(op, self._h, other._h if other else None))

    def intersect(self, o): return self._op(0, o)
    def union(self, *others):
        if len(others) == 1:
            return self._op(1, others[0])
        arr = (C.c_void_p * len(others))(*[o._h for o in others])
        return Bitmap(_h=lib().orc_bitmap_union_many(self._h, arr, len(others)))
    def difference(self, o): return self._op(2, o)
    def xor(self, o): return self._op(3, o)
    def shift(self): return self._op(4, None)

    def intersection_count(self, o: "Bitmap") -> int:
        return lib().orc_bitmap_intersection_count(self._h, o._h)

    def slice(self) -> List[int]:
        n = lib().orc_bitmap_slice(self._h, None, 0)
        buf = (C.c_uint64 * max(1, n))()
        lib().orc_bitmap_slice(self._h, buf, n)
        return list(buf[:n])

    def containers(self) -> List[Tuple[int, int, int]]:
        n = lib().orc_bitmap_containers(self._h, None, None, None, 0)
        k, t, c = (C.c_uint64 * max(1, n))(), (C.c_int * max(1, n))(), (C.c_int * max(1, n))()
        lib().orc_bitmap_containers(self._h, k, t, c, n)
        return [(k[i], t[i], c[i]) for i in range(n)]

    def serialize(self, optimize: bool = True) -> bytes:
        n = lib().orc_bitmap_serialize(self._h, int(optimize), None, 0)
        buf = C.create_string_buffer(max(1, n))
        lib().orc_bitmap_serialize(self._h, int(optimize), buf, n)
        return buf.raw[:n]

    @staticmethod
    def deserialize(data: bytes) -> "Bitmap":
        h = lib().orc_bitmap_deserialize(data, len(data))
        if not h:
            raise OracleError(lib().orc_last_error().decode())
        return Bitmap(_h=h)


def encode_op(typ: int, values: Sequence[int]) -> bytes:
    n = lib().orc_encode_op(typ, u64_array(values), len(values), None, 0)
    buf = C.create_string_buffer(n)
    lib().orc_encode_op(typ, u64_array(values), len(values), buf, n)
    return buf.raw[:n]


class Row:
    def __init__(self, *cols: int, _h=None):
        self._h = _h if _h is not None else lib().orc_row_new(u64_array(cols), len(cols))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_row_free(self._h)
            self._h = None

    def _op(self, op, o): return Row(_h=lib().orc_row_op(op, self._h, o._h if o else None))
    def intersect(self, o): return self._op(0, o)
    def union(self, o): return self._op(1, o)
    def difference(self, o): return self._op(2, o)
    def xor(self, o): return self._op(3, o)
    def shift(self): return self._op(4, None)
    def merge(self, o): return self._op(5, o)
    def count(self): return lib().orc_row_count(self._h)
    def intersection_count(self, o): return lib().orc_row_intersection_count(self._h, o._h)

    def columns(self) -> List[int]:
        n = lib().orc_row_columns(self._h, None, 0)
        buf = (C.c_uint64 * max(1, n))()
        lib().orc_row_columns(self._h, buf, n)
        return list(buf[:n])


# ---------------------------------------------------------------------------------------
# Holder + executor
# ---------------------------------------------------------------------------------------
class Oracle:
    """CPU executor over oracle fragments; method names mirror the engine binding."""

    def __init__(self, workers: int = 0):
        self._h = lib().orc_new(workers)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_free(self._h)
            self._h = None

    def set_workers(self, w: int) -> None:
        lib().orc_set_workers(self._h, w)

    def family_create(self) -> int:
        return lib().orc_family_create(self._h)

    def frag_load(self, fam: int, shard: int, data) -> None:
        if hasattr(data, "ptr"):  # synth.FragBytes: pass the generator's buffer without a copy
            _check(lib().orc_frag_load(self._h, fam, shard, C.cast(C.c_void_p(data.ptr), C.c_char_p), data.size))
        else:
            _check(lib().orc_frag_load(self._h, fam, shard, data, len(data)))

    def count_batch(self, progs, shards: Sequence[int], fused: bool = False) -> List[int]:
        return [self.count(p, shards, fused) for p in progs]

    def set_bit(self, fam, shard, row, col) -> bool:
        return bool(lib().orc_frag_set_bit(self._h, fam, shard, row, col))

    def clear_bit(self, fam, shard, row, col) -> bool:
        return bool(lib().orc_frag_clear_bit(self._h, fam, shard, row, col))

    def set_value(self, fam, shard, col, bit_depth, value) -> bool:
        return bool(lib().orc_frag_set_value(self._h, fam, shard, col, bit_depth, value))

    def clear_value(self, fam, shard, col, bit_depth, value) -> bool:
        return bool(lib().orc_frag_clear_value(self._h, fam, shard, col, bit_depth, value))

    def value(self, fam, shard, col, bit_depth):
        out = C.c_int64()
        ok = lib().orc_frag_value(self._h, fam, shard, col, bit_depth, C.byref(out))
        return (out.value, True) if ok else (0, False)

    def cache_config(self, fam, shard, cache_none=False, max_entries=50000) -> None:
        lib().orc_frag_cache_config(self._h, fam, shard, int(cache_none), max_entries)

    def cache_add(self, fam, shard, row_id: int, n: int) -> None:
        lib().orc_frag_cache_add(self._h, fam, shard, row_id, n)

    def cache_len(self, fam, shard) -> int:
        return lib().orc_frag_cache_len(self._h, fam, shard)

    def recalculate_cache(self, fam, shard) -> None:
        lib().orc_frag_recalculate_cache(self._h, fam, shard)

    def rebuild_cache(self, fam, shard) -> None:
        lib().orc_frag_rebuild_cache(self._h, fam, shard)

    def rankings(self, fam, shard) -> Tuple[List[int], List[int]]:
        n = lib().orc_frag_rankings(self._h, fam, shard, None, None, 0)
        ids, cnt = (C.c_uint64 * max(1, n))(), (C.c_uint64 * max(1, n))()
        lib().orc_frag_rankings(self._h, fam, shard, ids, cnt, n)
        return list(ids[:n]), list(cnt[:n])

    def frag_serialize(self, fam, shard) -> bytes:
        n = lib().orc_frag_serialize(self._h, fam, shard, None, 0)
        buf = C.create_string_buffer(max(1, n))
        lib().orc_frag_serialize(self._h, fam, shard, buf, n)
        return buf.raw[:n]

    def count(self, prog, shards: Sequence[int], fused: bool = False) -> int:
        out = C.c_uint64()
        _check(lib().orc_count(self._h, prog, len(prog), u64_array(shards), len(shards), int(fused), C.byref(out)))
        return out.value

    def count_shards(self, prog, shards, fused=False) -> List[int]:
        out = (C.c_uint64 * max(1, len(shards)))()
        _check(lib().orc_count_shards(self._h, prog, len(prog), u64_array(shards), len(shards), int(fused), out))
        return list(out[: len(shards)])

    def count_batch_raw(self, flat, off, n_queries: int, shards_arr, n_shards: int, fused=False, out=None):
        """n_queries Count programs (flat pb200_node array + offsets) in flight together on the shared worker pool."""
        if out is None:
            out = (C.c_uint64 * max(1, n_queries))()
        _check(lib().orc_count_batch(self._h, flat, off, n_queries, shards_arr, n_shards, int(fused), out))
        return out

    def eval_columns(self, prog, shard: int) -> List[int]:
        n = C.c_size_t()
        _check(lib().orc_eval_columns(self._h, prog, len(prog), shard, None, 0, C.byref(n)))
        buf = (C.c_uint64 * max(1, n.value))()
        _check(lib().orc_eval_columns(self._h, prog, len(prog), shard, buf, n.value, C.byref(n)))
        return list(buf[: n.value])

    def eval_roaring(self, prog, shard: int) -> bytes:
        n = C.c_size_t()
        _check(lib().orc_eval_roaring(self._h, prog, len(prog), shard, None, 0, C.byref(n)))
        buf = C.create_string_buffer(max(1, n.value))
        _check(lib().orc_eval_roaring(self._h, prog, len(prog), shard, buf, n.value, C.byref(n)))
        return buf.raw[: n.value]

    def eval_proto_columns(self, prog, shards) -> bytes:
        """The Columns field of internal.Row for the call's result over `shards` (public.pb.go:745-766)."""
        n = C.c_size_t()
        sh = u64_array(shards)
        _check(lib().orc_eval_proto_columns(self._h, prog, len(prog), sh, len(shards), None, 0, C.byref(n)))
        buf = C.create_string_buffer(max(1, n.value))
        _check(lib().orc_eval_proto_columns(self._h, prog, len(prog), sh, len(shards), buf, n.value, C.byref(n)))
        return buf.raw[: n.value]

    def eval_roaring_shards(self, prog, shards) -> list:
        return [self.eval_roaring(prog, s) for s in shards]

    def _pairs(self, fn, *args, cap=1 << 16):
        out = (Pair * cap)()
        n = C.c_size_t()
        _check(fn(*args, out, cap, C.byref(n)))
        return [(out[i].id, out[i].count) for i in range(min(cap, n.value))]

    def topn(self, fam, src_prog, n, shards, ids=(), min_threshold=0, tanimoto=0, remote=False):
        sp, ns = (src_prog, len(src_prog)) if src_prog is not None else (None, 0)
        return self._pairs(lib().orc_topn, self._h, fam, sp, ns, n, u64_array(ids) if ids else None, len(ids),
                           min_threshold, tanimoto, 1 if remote else 0, u64_array(shards), len(shards))

    def topn_shard(self, fam, src_prog, n, shard, ids=(), min_threshold=0, tanimoto=0):
        sp, ns = (src_prog, len(src_prog)) if src_prog is not None else (None, 0)
        return self._pairs(lib().orc_topn_shard, self._h, fam, sp, ns, n, u64_array(ids) if ids else None, len(ids),
                           min_threshold, tanimoto, shard)

    def frag_drop(self, fam, shard) -> None:
        _check(lib().orc_frag_drop(self._h, fam, shard))

    def frag_import_positions(self, fam, shard, set_pos=(), clear_pos=()) -> int:
        """fragment.importPositions (fragment.go:2051-2098): AddN / RemoveN + the rank-cache recount of the rows named."""
        n = C.c_int64()
        _check(lib().orc_frag_import_positions(self._h, fam, shard, u64_array(set_pos), len(set_pos), u64_array(clear_pos), len(clear_pos), C.byref(n)))
        return n.value

    def frag_import_roaring(self, fam, shard, data: bytes, clear: bool = False) -> int:
        n = C.c_int64()
        _check(lib().orc_frag_import_roaring(self._h, fam, shard, data, len(data), int(clear), C.byref(n)))
        return n.value

    def rows(self, fam, shards, start=0, column=None, limit=None):
        from pilosa_b200.abi import NO_LIMIT, ROWS_COLUMN
        cap = 1 << 16
        while True:
            out, n = (C.c_uint64 * cap)(), C.c_size_t()
            _check(lib().orc_rows(self._h, fam, start, 0 if column is None else ROWS_COLUMN, column or 0,
                                  NO_LIMIT if limit is None else limit, u64_array(shards), len(shards), out, cap, C.byref(n)))
            if n.value <= cap:
                return list(out[: n.value])
            cap = n.value

    def groupby(self, children, filter_prog, limit, shards):
        """children: [(family, previous or None, rows or None)]; returns [((row ids), count)]."""
        from pilosa_b200.abi import NO_LIMIT, Group, pack_groupby_children, unpack_groups
        arr, keep = pack_groupby_children(children)
        fp, nf = (filter_prog, len(filter_prog)) if filter_prog is not None else (None, 0)
        cap = 1 << 14
        while True:
            out, n = (Group * cap)(), C.c_size_t()
            _check(lib().orc_groupby(self._h, arr, len(children), fp, nf, NO_LIMIT if limit is None else limit,
                                     u64_array(shards), len(shards), out, cap, C.byref(n)))
            if n.value <= cap:
                return unpack_groups(out, n.value, len(children))
            cap = n.value

    def bsi_sum(self, fam, bit_depth, filter_prog, shards, flags=0):
        fp, nf = (filter_prog, len(filter_prog)) if filter_prog is not None else (None, 0)
        s, c = C.c_int64(), C.c_uint64()
        _check(lib().orc_bsi_sum(self._h, fam, bit_depth, fp, nf, flags, u64_array(shards), len(shards), C.byref(s), C.byref(c)))
        return s.value, c.value

    def bsi_minmax(self, fam, bit_depth, filter_prog, shards, is_max):
        fp, nf = (filter_prog, len(filter_prog)) if filter_prog is not None else (None, 0)
        v, c = C.c_int64(), C.c_uint64()
        _check(lib().orc_bsi_minmax(self._h, fam, bit_depth, fp, nf, int(is_max), u64_array(shards), len(shards), C.byref(v), C.byref(c)))
        return v.value, c.value

    def bsi_min(self, fam, bit_depth, filter_prog, shards): return self.bsi_minmax(fam, bit_depth, filter_prog, shards, False)
    def bsi_max(self, fam, bit_depth, filter_prog, shards): return self.bsi_minmax(fam, bit_depth, filter_prog, shards, True)

    def bsi_minmax_shards(self, fam, bit_depth, filter_prog, shards, is_max):
        fp, nf = (filter_prog, len(filter_prog)) if filter_prog is not None else (None, 0)
        v, c = (C.c_int64 * max(1, len(shards)))(), (C.c_uint64 * max(1, len(shards)))()
        _check(lib().orc_bsi_minmax_shards(self._h, fam, bit_depth, fp, nf, int(is_max), u64_array(shards), len(shards), v, c))
        return list(v[: len(shards)]), list(c[: len(shards)])


def base_value(base: int, bit_depth: int, op: int, value: int) -> Tuple[int, bool]:
    oor = C.c_int()
    v = lib().orc_base_value(base, bit_depth, op, value, C.byref(oor))
    return v, bool(oor.value)


def base_value_between(base: int, bit_depth: int, lo: int, hi: int) -> Tuple[int, int, bool]:
    olo, ohi = C.c_int64(), C.c_int64()
    oor = lib().orc_base_value_between(base, bit_depth, lo, hi, C.byref(olo), C.byref(ohi))
    return olo.value, ohi.value, bool(oor)
