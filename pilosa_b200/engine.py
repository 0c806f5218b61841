"""ctypes binding of libpilosa_b200.so (include/pilosa_b200.h).

There is no fallback of any kind: if the shared library is missing or a declared symbol is
not exported, ``load_library`` raises; if no CUDA device is usable ``Engine()`` raises.
"""
from __future__ import annotations

import ctypes as C
import os
import re
from typing import List, Optional, Sequence, Tuple

from .abi import Node, Pair, Timing, u64_array

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpilosa_b200.so")
HEADER_PATH = os.path.join(_HERE, "..", "include", "pilosa_b200.h")


class EngineError(RuntimeError):
    def __init__(self, status: int, msg: str):
        super().__init__(f"pb200 status {status}: {msg}")
        self.status = status


def declared_symbols() -> List[str]:
    """Every function the public header declares (used by the ABI-completeness test)."""
    src = open(HEADER_PATH).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(pb200_[a-z0-9_]+)\s*\(", src)))


_lib: Optional[C.CDLL] = None


def load_library() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python __graft_entry__.py` "
            "(nvcc -gencode arch=compute_100a,code=sm_100a). pilosa_b200 has no CPU fallback."
        )
    L = C.CDLL(LIB_PATH)
    missing = [s for s in declared_symbols() if not hasattr(L, s)]
    if missing:
        raise ImportError(f"{LIB_PATH} does not export: {missing}")
    vp, sz = C.c_void_p, C.c_size_t
    u64p, i64p, u32p = C.POINTER(C.c_uint64), C.POINTER(C.c_int64), C.POINTER(C.c_uint32)
    np_, pp = C.POINTER(Node), C.POINTER(Pair)
    u32, u64, i32 = C.c_uint32, C.c_uint64, C.c_int
    sig = {
        "pb200_init": (i32, [i32, sz, C.POINTER(vp)]),
        "pb200_shutdown": (i32, [vp]),
        "pb200_strerror": (C.c_char_p, [i32]),
        "pb200_last_error": (C.c_char_p, [vp]),
        "pb200_abi_version": (i32, []),
        "pb200_device_count": (i32, []),
        "pb200_set_stream": (i32, [vp, vp]),
        "pb200_family_create": (i32, [vp, u32p]),
        "pb200_frag_load": (i32, [vp, u32, u64, C.c_char_p, sz]),
        "pb200_frag_apply": (i32, [vp, u32, u64, u64p, sz, u64p, sz]),
        "pb200_frag_drop": (i32, [vp, u32, u64]),
        "pb200_rank_set": (i32, [vp, u32, u64, u64p, u64p, sz]),
        "pb200_rank_get": (i32, [vp, u32, u64, u64p, u64p, sz, C.POINTER(sz)]),
        "pb200_host_check_roaring": (i32, [C.c_char_p, sz, u64p, u64p, C.c_char_p, sz]),
        "pb200_frag_stat": (i32, [vp, u32, u64, u64p, u64p, u64p]),
        "pb200_count": (i32, [vp, np_, u32, u64p, u32, u64p]),
        "pb200_count_batch": (i32, [vp, np_, u32p, u32, u64p, u32, u64p]),
        "pb200_count_shards": (i32, [vp, np_, u32, u64p, u32, u64p]),
        "pb200_count_batch_begin": (i32, [vp, np_, u32p, u32, u64p, u32]),
        "pb200_count_batch_device": (i32, [vp, C.POINTER(vp)]),
        "pb200_count_batch_end": (i32, [vp, u64p]),
        "pb200_batch_prepare": (i32, [vp, np_, u32p, u32, u64p, u32, C.POINTER(vp)]),
        "pb200_batch_launch": (i32, [vp, vp]),
        "pb200_batch_device_counts": (i32, [vp, vp, C.POINTER(vp)]),
        "pb200_batch_fetch": (i32, [vp, vp, u64p]),
        "pb200_batch_free": (i32, [vp, vp]),
        "pb200_count_pairs_noalloc": (i32, [vp, u32, vp, u32, vp, u32, vp]),
        "pb200_eval": (i32, [vp, np_, u32, u64, C.c_char_p, sz, C.POINTER(sz)]),
        "pb200_eval_columns": (i32, [vp, np_, u32, u64, u64p, sz, C.POINTER(sz)]),
        "pb200_frag_import_roaring": (i32, [vp, u32, u64, C.c_char_p, sz, i32, C.POINTER(C.c_int64)]),
        "pb200_set_option": (i32, [vp, u32, u64]),
        "pb200_host_register": (i32, [vp, vp, sz]),
        "pb200_host_unregister": (i32, [vp, vp]),
        "pb200_rows": (i32, [vp, u32, u64, u32, u64, u64, u64p, u32, u64p, sz, C.POINTER(sz)]),
        "pb200_groupby": (i32, [vp, vp, u32, np_, u32, u64, u64p, u32, vp, sz, C.POINTER(sz)]),
        "pb200_eval_shards": (i32, [vp, np_, u32, u64p, u32, vp, sz, u64p, C.POINTER(sz)]),
        "pb200_eval_proto_columns": (i32, [vp, np_, u32, u64p, u32, vp, sz, C.POINTER(sz)]),
        "pb200_topn": (i32, [vp, u32, np_, u32, u32, u64p, sz, u64, u64, u32, u64p, u32, pp, sz, C.POINTER(sz)]),
        "pb200_topn_shard": (i32, [vp, u32, np_, u32, u32, u64p, sz, u64, u64, u64, pp, sz, C.POINTER(sz)]),
        "pb200_bsi_sum": (i32, [vp, u32, u32, np_, u32, u32, u64p, u32, i64p, u64p]),
        "pb200_bsi_min": (i32, [vp, u32, u32, np_, u32, u64p, u32, i64p, u64p]),
        "pb200_bsi_max": (i32, [vp, u32, u32, np_, u32, u64p, u32, i64p, u64p]),
        "pb200_bsi_minmax_shards": (i32, [vp, u32, u32, np_, u32, i32, u64p, u32, i64p, u64p]),
        "pb200_last_timing": (i32, [vp, C.POINTER(Timing)]),
        "pb200_reserve": (i32, [vp, u32, u32, u32]),
        "pb200_combine_stats": (i32, [vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
        "pb200_peer_export": (i32, [vp, C.c_char_p]),
        "pb200_peer_connect": (i32, [vp, i32, i32, C.c_char_p]),
        "pb200_peer_flush": (i32, [vp]),
        "pb200_batch_launch_allreduce": (i32, [vp, vp]),
    }
    # pb200_multi_*: the same argument lists behind the multi-GPU handle
    for name in MULTI_FORWARDED:
        sig["pb200_multi_" + name] = sig["pb200_" + name]
    sig["pb200_init_multi"] = (i32, [C.POINTER(i32), i32, sz, C.POINTER(vp)])
    sig["pb200_multi_shutdown"] = (i32, [vp])
    sig["pb200_multi_last_error"] = (C.c_char_p, [vp])
    sig["pb200_multi_ndev"] = (i32, [vp])
    sig["pb200_multi_peer_access"] = (i32, [vp])
    sig["pb200_multi_ctx"] = (i32, [vp, i32, C.POINTER(vp)])
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype, fn.argtypes = res, args
    _lib = L
    return L


# single-device entry points that exist, with the same arguments, as pb200_multi_<name>
MULTI_FORWARDED = [
    "family_create", "frag_load", "frag_apply", "frag_import_roaring", "frag_drop", "rank_set", "rank_get", "frag_stat",
    "count", "count_batch", "count_shards", "eval", "eval_columns", "eval_shards", "eval_proto_columns", "rows", "groupby",
    "topn", "topn_shard", "bsi_sum", "bsi_min", "bsi_max", "bsi_minmax_shards", "last_timing", "set_option",
    "host_register", "host_unregister",
]


class _MultiLib:
    """Presents pb200_multi_<name> under the name pb200_<name>, so that Engine's methods drive the multi-GPU handle."""

    def __init__(self, lib):
        self._lib = lib

    def __getattr__(self, name):
        if name in ("pb200_strerror", "pb200_abi_version", "pb200_host_check_roaring"):
            return getattr(self._lib, name)
        if name == "pb200_last_error":
            return self._lib.pb200_multi_last_error
        if name == "pb200_shutdown":
            return self._lib.pb200_multi_shutdown
        if name.startswith("pb200_") and name[6:] in MULTI_FORWARDED:
            return getattr(self._lib, "pb200_multi_" + name[6:])
        raise AttributeError(f"{name} has no multi-GPU form")


def device_count() -> int:
    return int(load_library().pb200_device_count())


def host_check_roaring(data: bytes) -> Tuple[int, int]:
    """Validate fragment-file bytes with the loader's parser, no GPU needed.
    Returns (non-empty containers, total bits) or raises EngineError(PB200_ERR_FORMAT)."""
    lib = load_library()
    nc, nb = C.c_uint64(), C.c_uint64()
    err = C.create_string_buffer(256)
    rc = lib.pb200_host_check_roaring(data, len(data), C.byref(nc), C.byref(nb), err, 256)
    if rc != 0:
        raise EngineError(rc, err.value.decode() or lib.pb200_strerror(rc).decode())
    return nc.value, nb.value


class Engine:
    """One context bound to one CUDA device (pb200_ctx)."""

    def __init__(self, device: int = 0, hbm_budget_bytes: int = 0):
        self.lib = load_library()
        h = C.c_void_p()
        rc = self.lib.pb200_init(device, hbm_budget_bytes, C.byref(h))
        if rc != 0:
            raise EngineError(rc, f"pb200_init(device={device}) failed: {self.lib.pb200_strerror(rc).decode()} "
                                  "(a CUDA device is required; there is no CPU fallback)")
        self._h = h
        self._rows_buf = None
        self.device = device

    OPT_DENSIFY_MIN_BYTES = 1
    OPT_RANK_CACHE_SIZE = 2

    def set_option(self, option: int, value: int) -> None:
        self._check(self.lib.pb200_set_option(self._h, option, value))

    def _release_rows_buf(self) -> None:
        if getattr(self, "_rows_pinned", False) and getattr(self, "_h", None):
            self.lib.pb200_host_unregister(self._h, C.addressof(self._rows_c))
        self._rows_pinned = False
        self._rows_c = None
        self._rows_buf = None

    def close(self) -> None:
        if getattr(self, "_h", None):
            self._release_rows_buf()
            self._release_proto_buf()
            self.lib.pb200_shutdown(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc: int) -> None:
        if rc != 0:
            raise EngineError(rc, f"{self.lib.pb200_strerror(rc).decode()}: {self.lib.pb200_last_error(self._h).decode()}")

    # ---- fragments ----
    def set_stream(self, cuda_stream: int) -> None:
        self._check(self.lib.pb200_set_stream(self._h, C.c_void_p(cuda_stream)))

    def family_create(self) -> int:
        out = C.c_uint32()
        self._check(self.lib.pb200_family_create(self._h, C.byref(out)))
        return out.value

    def frag_load(self, fam: int, shard: int, data) -> None:
        """data: bytes, or any object with .ptr/.size (synth.FragBytes) passed without a copy."""
        if hasattr(data, "ptr"):
            self._check(self.lib.pb200_frag_load(self._h, fam, shard, C.cast(C.c_void_p(data.ptr), C.c_char_p), data.size))
        else:
            self._check(self.lib.pb200_frag_load(self._h, fam, shard, data, len(data)))

    def frag_apply(self, fam: int, shard: int, set_pos: Sequence[int] = (), clear_pos: Sequence[int] = ()) -> None:
        self._check(self.lib.pb200_frag_apply(self._h, fam, shard, u64_array(set_pos), len(set_pos), u64_array(clear_pos), len(clear_pos)))

    def frag_import_roaring(self, fam: int, shard: int, data: bytes, clear: bool = False) -> int:
        """fragment.importRoaring: union / subtract a serialized bitmap (Pilosa or official format); returns bits changed."""
        n = C.c_int64()
        self._check(self.lib.pb200_frag_import_roaring(self._h, fam, shard, data, len(data), int(clear), C.byref(n)))
        return n.value

    def frag_drop(self, fam: int, shard: int) -> None:
        self._check(self.lib.pb200_frag_drop(self._h, fam, shard))

    def rank_get(self, fam: int, shard: int) -> Tuple[List[int], List[int]]:
        """The fragment's TopN candidate order (rankCache.Top(), cache.go:288): (row ids, cached counts)."""
        n = C.c_size_t()
        self._check(self.lib.pb200_rank_get(self._h, fam, shard, None, None, 0, C.byref(n)))
        ids, cnt = (C.c_uint64 * max(1, n.value))(), (C.c_uint64 * max(1, n.value))()
        self._check(self.lib.pb200_rank_get(self._h, fam, shard, ids, cnt, n.value, C.byref(n)))
        return list(ids[: n.value]), list(cnt[: n.value])

    def rank_set(self, fam: int, shard: int, row_ids: Sequence[int], counts: Sequence[int]) -> None:
        assert len(row_ids) == len(counts)
        self._check(self.lib.pb200_rank_set(self._h, fam, shard, u64_array(row_ids), u64_array(counts), len(row_ids)))

    def frag_stat(self, fam: int, shard: int) -> Tuple[int, int, int]:
        a, b, c = C.c_uint64(), C.c_uint64(), C.c_uint64()
        self._check(self.lib.pb200_frag_stat(self._h, fam, shard, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    # ---- queries ----
    def count(self, prog, shards: Sequence[int]) -> int:
        out = C.c_uint64()
        self._check(self.lib.pb200_count(self._h, prog, len(prog), u64_array(shards), len(shards), C.byref(out)))
        return out.value

    def count_shards(self, prog, shards: Sequence[int]) -> List[int]:
        out = (C.c_uint64 * max(1, len(shards)))()
        self._check(self.lib.pb200_count_shards(self._h, prog, len(prog), u64_array(shards), len(shards), out))
        return list(out[: len(shards)])

    def count_batch(self, progs: Sequence, shards: Sequence[int]) -> List[int]:
        """progs: list of Node arrays."""
        total = sum(len(p) for p in progs)
        flat = (Node * max(1, total))()
        off = (C.c_uint32 * (len(progs) + 1))()
        k = 0
        for i, p in enumerate(progs):
            off[i] = k
            for nd in p:
                flat[k] = nd
                k += 1
        off[len(progs)] = k
        return self.count_batch_raw(flat, off, len(progs), u64_array(shards), len(shards))

    def count_batch_raw(self, flat, off, n_queries: int, shards_arr, n_shards: int, out=None) -> List[int]:
        """Zero-marshalling variant for the bench: ctypes arrays prepared once by the caller."""
        out = out if out is not None else (C.c_uint64 * max(1, n_queries))()
        self._check(self.lib.pb200_count_batch(self._h, flat, off, n_queries, shards_arr, n_shards, out))
        return out

    # ---- split / prepared Count batches (multi-GPU reduce, device-resident serving loop) ----
    @staticmethod
    def pack_programs(progs: Sequence):
        """Concatenate Node arrays into the (flat, prog_off) pair the batch entry points take."""
        total = sum(len(p) for p in progs)
        flat = (Node * max(1, total))()
        off = (C.c_uint32 * (len(progs) + 1))()
        k = 0
        for i, p in enumerate(progs):
            off[i] = k
            for nd in p:
                flat[k] = nd
                k += 1
        off[len(progs)] = k
        return flat, off

    def count_batch_begin(self, flat, off, n_queries: int, shards_arr, n_shards: int) -> None:
        self._check(self.lib.pb200_count_batch_begin(self._h, flat, off, n_queries, shards_arr, n_shards))

    def count_batch_device(self) -> int:
        p = C.c_void_p()
        self._check(self.lib.pb200_count_batch_device(self._h, C.byref(p)))
        return p.value or 0

    def count_batch_end(self, out) -> None:
        self._check(self.lib.pb200_count_batch_end(self._h, out))

    def batch_prepare(self, flat, off, n_queries: int, shards_arr, n_shards: int):
        b = C.c_void_p()
        self._check(self.lib.pb200_batch_prepare(self._h, flat, off, n_queries, shards_arr, n_shards, C.byref(b)))
        return b

    def batch_launch(self, b) -> None:
        self._check(self.lib.pb200_batch_launch(self._h, b))

    # ---- one process per GPU: counts summed over the ranks by the batch's own kernel (NVLink peer memory, no NCCL) ----
    OPT_PEER_ALLREDUCE = 4
    PEER_HANDLE_BYTES = 64

    def reserve(self, max_queries: int, max_shards: int, max_nodes_per_query: int = 3) -> None:
        """Pre-size the scratch of the Count paths so that serving never allocates (pb200_reserve)."""
        self._check(self.lib.pb200_reserve(self._h, max_queries, max_shards, max_nodes_per_query))

    def combine_stats(self):
        """(combined launches, single-query calls they served) of pb200_count's combining queue."""
        a, b = C.c_uint64(), C.c_uint64()
        self._check(self.lib.pb200_combine_stats(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def peer_export(self) -> bytes:
        buf = C.create_string_buffer(self.PEER_HANDLE_BYTES)
        self._check(self.lib.pb200_peer_export(self._h, buf))
        return buf.raw

    def peer_connect(self, rank: int, world: int, handles: bytes) -> None:
        assert len(handles) == world * self.PEER_HANDLE_BYTES
        self._check(self.lib.pb200_peer_connect(self._h, rank, world, handles))

    def peer_flush(self) -> None:
        self._check(self.lib.pb200_peer_flush(self._h))

    def batch_launch_allreduce(self, b) -> None:
        self._check(self.lib.pb200_batch_launch_allreduce(self._h, b))

    def batch_device_counts(self, b) -> int:
        p = C.c_void_p()
        self._check(self.lib.pb200_batch_device_counts(self._h, b, C.byref(p)))
        return p.value or 0

    def batch_fetch(self, b, out) -> None:
        self._check(self.lib.pb200_batch_fetch(self._h, b, out))

    def batch_free(self, b) -> None:
        self._check(self.lib.pb200_batch_free(self._h, b))

    def count_pairs_noalloc(self, fam: int, d_pairs: int, n_pairs: int, d_shards: int, n_shards: int, d_counts: int) -> None:
        """Allocation-free Count(Intersect) over row pairs. All three buffers are DEVICE
        addresses owned by the caller (e.g. ``tensor.data_ptr()``); the call only enqueues work
        on the context's stream — synchronise that stream before reading ``d_counts``."""
        self._check(self.lib.pb200_count_pairs_noalloc(self._h, fam, C.c_void_p(d_pairs), n_pairs, C.c_void_p(d_shards), n_shards,
                                                       C.c_void_p(d_counts)))

    def eval_roaring(self, prog, shard: int) -> bytes:
        n = C.c_size_t()
        self._check(self.lib.pb200_eval(self._h, prog, len(prog), shard, None, 0, C.byref(n)))
        buf = C.create_string_buffer(max(1, n.value))
        self._check(self.lib.pb200_eval(self._h, prog, len(prog), shard, buf, n.value, C.byref(n)))
        return buf.raw[: n.value]

    def eval_shards_raw(self, prog, shards: Sequence[int]):
        """pb200_eval_shards into a reused host buffer: (buffer, offsets[n_shards + 1])."""
        ns = len(shards)
        cap = max(1, ns) * (131072 + 536)
        if self._rows_buf is None or len(self._rows_buf) < cap:
            self._release_rows_buf()
            self._rows_buf = bytearray(cap)
            self._rows_c = (C.c_char * cap).from_buffer(self._rows_buf)
            # page-locked once: the packed streams then arrive by direct DMA (pb200_host_register)
            self._rows_pinned = self.lib.pb200_host_register(self._h, C.addressof(self._rows_c), cap) == 0
        buf = self._rows_c
        off = (C.c_uint64 * (ns + 1))()
        n = C.c_size_t()
        self._check(self.lib.pb200_eval_shards(self._h, prog, len(prog), u64_array(shards) if ns else None, ns, buf, cap, off, C.byref(n)))
        return self._rows_buf, off

    def _proto_alloc(self, cap: int) -> None:
        self._release_proto_buf()
        self._proto_buf = C.create_string_buffer(cap)
        self._proto_pinned = self.lib.pb200_host_register(self._h, C.addressof(self._proto_buf), cap) == 0

    def _release_proto_buf(self) -> None:
        if getattr(self, "_proto_pinned", False) and getattr(self, "_h", None):
            self.lib.pb200_host_unregister(self._h, C.addressof(self._proto_buf))
        self._proto_pinned = False
        self._proto_buf = None

    def eval_proto_columns_raw(self, prog, shards: Sequence[int]):
        """pb200_eval_proto_columns into a reused, page-locked host buffer: (buffer, length). One call when the
        buffer is large enough; PB200_ERR_RANGE reports the size to grow it to."""
        ns = len(shards)
        sh = u64_array(shards) if ns else None
        n = C.c_size_t()
        if getattr(self, "_proto_buf", None) is None:
            self._proto_alloc(1 << 20)
        rc = self.lib.pb200_eval_proto_columns(self._h, prog, len(prog), sh, ns, self._proto_buf, len(self._proto_buf), C.byref(n))
        if rc == -6:  # PB200_ERR_RANGE
            self._proto_alloc(n.value + n.value // 4)
            rc = self.lib.pb200_eval_proto_columns(self._h, prog, len(prog), sh, ns, self._proto_buf, len(self._proto_buf), C.byref(n))
        self._check(rc)
        return self._proto_buf, n.value

    def eval_proto_columns(self, prog, shards: Sequence[int]) -> bytes:
        """The Columns field of internal.Row for the call's result over `shards`, encoded on the device."""
        buf, n = self.eval_proto_columns_raw(prog, shards)
        return C.string_at(buf, n)

    def eval_roaring_shards(self, prog, shards: Sequence[int]) -> List[bytes]:
        """One row segment (Pilosa roaring bytes) per shard, typed and serialised on the device."""
        if not shards:
            return []
        buf, off = self.eval_shards_raw(prog, shards)
        mv = memoryview(buf)
        return [bytes(mv[off[i]: off[i + 1]]) for i in range(len(shards))]

    def rows(self, fam, shards, start=0, column=None, limit=None) -> List[int]:
        from .abi import NO_LIMIT, ROWS_COLUMN
        cap = 1 << 10
        while True:
            out, n = (C.c_uint64 * cap)(), C.c_size_t()
            rc = self.lib.pb200_rows(self._h, fam, start, 0 if column is None else ROWS_COLUMN, column or 0,
                                     NO_LIMIT if limit is None else limit, u64_array(shards), len(shards), out, cap, C.byref(n))
            if rc == -6 and n.value > cap:  # PB200_ERR_RANGE: retry with the size the call reported
                cap = n.value
                continue
            self._check(rc)
            return list(out[: n.value])

    def groupby(self, children, filter_prog, limit, shards):
        """children: [(family, previous or None, rows or None)]; returns [((row ids), count)]."""
        from .abi import NO_LIMIT, Group, pack_groupby_children, unpack_groups
        arr, keep = pack_groupby_children(children)
        fp, nf = (filter_prog, len(filter_prog)) if filter_prog is not None else (None, 0)
        cap = getattr(self, "_gb_cap", 1 << 10)
        while True:
            out, n = (Group * cap)(), C.c_size_t()
            rc = self.lib.pb200_groupby(self._h, arr, len(children), fp, nf, NO_LIMIT if limit is None else limit,
                                        u64_array(shards), len(shards), out, cap, C.byref(n))
            if rc == -6 and n.value > cap:
                cap = self._gb_cap = n.value  # remembered: the retry runs the query again
                continue
            self._check(rc)
            return unpack_groups(out, n.value, len(children))

    def eval_columns(self, prog, shard: int) -> List[int]:
        n = C.c_size_t()
        self._check(self.lib.pb200_eval_columns(self._h, prog, len(prog), shard, None, 0, C.byref(n)))
        buf = (C.c_uint64 * max(1, n.value))()
        self._check(self.lib.pb200_eval_columns(self._h, prog, len(prog), shard, buf, n.value, C.byref(n)))
        return list(buf[: n.value])

    def _pairs(self, fn, *args, cap=1 << 10):
        while True:  # small buffer first (allocating a big ctypes array costs more than a TopN call)
            out = (Pair * cap)()
            n = C.c_size_t()
            rc = fn(*args, out, cap, C.byref(n))
            if rc == -6 and n.value > cap:  # PB200_ERR_RANGE: *n_out holds the size needed
                cap = n.value
                continue
            self._check(rc)
            return [(out[i].id, out[i].count) for i in range(n.value)]

    def topn(self, fam, src_prog, n, shards, ids=(), min_threshold=0, tanimoto=0, remote=False):
        sp, ns = (src_prog, len(src_prog)) if src_prog is not None else (None, 0)
        return self._pairs(self.lib.pb200_topn, self._h, fam, sp, ns, n, u64_array(ids) if ids else None, len(ids),
                           min_threshold, tanimoto, 1 if remote else 0, u64_array(shards), len(shards))

    def topn_shard(self, fam, src_prog, n, shard, ids=(), min_threshold=0, tanimoto=0):
        sp, ns = (src_prog, len(src_prog)) if src_prog is not None else (None, 0)
        return self._pairs(self.lib.pb200_topn_shard, self._h, fam, sp, ns, n, u64_array(ids) if ids else None, len(ids),
                           min_threshold, tanimoto, shard)

    def bsi_sum(self, fam, bit_depth, filter_prog, shards, flags=0):
        fp, nf = (filter_prog, len(filter_prog)) if filter_prog is not None else (None, 0)
        s, c = C.c_int64(), C.c_uint64()
        self._check(self.lib.pb200_bsi_sum(self._h, fam, bit_depth, fp, nf, flags, u64_array(shards), len(shards), C.byref(s), C.byref(c)))
        return s.value, c.value

    def bsi_min(self, fam, bit_depth, filter_prog, shards):
        fp, nf = (filter_prog, len(filter_prog)) if filter_prog is not None else (None, 0)
        v, c = C.c_int64(), C.c_uint64()
        self._check(self.lib.pb200_bsi_min(self._h, fam, bit_depth, fp, nf, u64_array(shards), len(shards), C.byref(v), C.byref(c)))
        return v.value, c.value

    def bsi_max(self, fam, bit_depth, filter_prog, shards):
        fp, nf = (filter_prog, len(filter_prog)) if filter_prog is not None else (None, 0)
        v, c = C.c_int64(), C.c_uint64()
        self._check(self.lib.pb200_bsi_max(self._h, fam, bit_depth, fp, nf, u64_array(shards), len(shards), C.byref(v), C.byref(c)))
        return v.value, c.value

    def bsi_minmax_shards(self, fam, bit_depth, filter_prog, shards, is_max):
        fp, nf = (filter_prog, len(filter_prog)) if filter_prog is not None else (None, 0)
        v, c = (C.c_int64 * max(1, len(shards)))(), (C.c_uint64 * max(1, len(shards)))()
        self._check(self.lib.pb200_bsi_minmax_shards(self._h, fam, bit_depth, fp, nf, int(is_max), u64_array(shards), len(shards), v, c))
        return list(v[: len(shards)]), list(c[: len(shards)])

    def last_timing(self) -> Timing:
        t = Timing()
        self._check(self.lib.pb200_last_timing(self._h, C.byref(t)))
        return t


class MultiEngine(Engine):
    """pb200_multi: the GPUs of one box behind one handle (one process). Same methods as Engine; every query takes the
    whole shard list — partition (shard % ndev), fan-out and reduce happen inside the library call."""

    OPT_SMALL_CALL_BYTES = 100

    def __init__(self, devices: Sequence[int], hbm_budget_bytes_per_device: int = 0):
        raw = load_library()
        self.lib = _MultiLib(raw)
        self._raw = raw
        h = C.c_void_p()
        arr = (C.c_int * len(devices))(*devices)
        rc = raw.pb200_init_multi(arr, len(devices), hbm_budget_bytes_per_device, C.byref(h))
        if rc != 0:
            raise EngineError(rc, f"pb200_init_multi(devices={list(devices)}) failed: {raw.pb200_strerror(rc).decode()} "
                                  "(CUDA devices are required; there is no CPU fallback)")
        self._h = h
        self._rows_buf = None
        self.devices = list(devices)

    @property
    def peer_access(self) -> bool:
        return bool(self._raw.pb200_multi_peer_access(self._h))

    def device_timing(self, i: int) -> Timing:
        c, t = C.c_void_p(), Timing()
        self._check(self._raw.pb200_multi_ctx(self._h, i, C.byref(c)))
        self._raw.pb200_last_timing(c, C.byref(t))
        return t
