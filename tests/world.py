"""Test helper: a tiny 'holder' built bit by bit in the CPU oracle (like the reference's tests
build fragments with setBit / setValue) and optionally mirrored into the CUDA engine by
handing it the oracle's serialized fragment files plus the oracle's rank-cache order
(pb200_frag_load / pb200_rank_set) — the same hand-over the Go host would do."""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import oracle
from pilosa_b200 import abi
from pilosa_b200.executor import Executor

SW = 1 << 20


class World:
    def __init__(self, backend: str):
        self.backend = backend
        self.orc = oracle.Oracle(workers=2)
        self.cpu = Executor(self.orc)
        self.eng = None
        self.gpu: Optional[Executor] = None
        if backend == "gpu":
            from pilosa_b200.engine import Engine

            self.eng = Engine(0)
            self.gpu = Executor(self.eng)
        self._dirty: set = set()           # (field, view, shard) changed since the last sync
        self._flt_rows = 0

    # ---- schema ----
    def create_field(self, name: str, type: str = "set", min: int = 0, max: int = 0, cache_size: int = 50000, cache_none: bool = False):
        for ex in self._execs():
            ex.create_field(name, type=type, min=min, max=max)
        self._cache = getattr(self, "_cache", {})
        self._cache[name] = (cache_none, cache_size)

    def _execs(self):
        return [self.cpu] + ([self.gpu] if self.gpu else [])

    def _view(self, field: str) -> str:
        f = self.cpu.fields[field]
        return "bsig_" + field if f.type == "int" else "standard"

    def _touch(self, field: str, shard: int):
        key = (field, self._view(field), shard)
        if key not in self._dirty and field in getattr(self, "_cache", {}):
            pass
        self._dirty.add(key)

    def _fam(self, field: str) -> int:
        return self.cpu.family(field, self._view(field))

    # ---- writes (always through the oracle: fragment.setBit / setValue semantics) ----
    def set_bit(self, field: str, row: int, col: int) -> bool:
        shard = col // SW
        fam = self._fam(field)
        self._ensure_cache(field, fam, shard)
        self._touch(field, shard)
        return self.orc.set_bit(fam, shard, row, col)

    def set_bits(self, field: str, row: int, *cols: int):  # mustSetBits
        for c in cols:
            self.set_bit(field, row, c)

    def clear_bit(self, field: str, row: int, col: int) -> bool:
        self._touch(field, col // SW)
        return self.orc.clear_bit(self._fam(field), col // SW, row, col)

    def set_value(self, field: str, col: int, value: int, bit_depth: Optional[int] = None) -> bool:
        bd = bit_depth if bit_depth is not None else self.cpu.fields[field].bsig.bit_depth
        self._touch(field, col // SW)
        return self.orc.set_value(self._fam(field), col // SW, col, bd, value)

    def clear_value(self, field: str, col: int, value: int, bit_depth: Optional[int] = None) -> bool:
        bd = bit_depth if bit_depth is not None else self.cpu.fields[field].bsig.bit_depth
        self._touch(field, col // SW)
        return self.orc.clear_value(self._fam(field), col // SW, col, bd, value)

    def _ensure_cache(self, field: str, fam: int, shard: int):
        cache_none, size = self._cache.get(field, (False, 50000))
        key = ("cfg", field, shard)
        if key not in self._dirty:
            self.orc.cache_config(fam, shard, cache_none, size)
            self._dirty.add(key)

    def recalculate_cache(self, field: str, shard: int = 0):
        self.orc.recalculate_cache(self._fam(field), shard)
        self._touch(field, shard)

    def new_row(self, *cols: int) -> abi.Call:
        """NewRow(cols...) of the reference's tests, as a leaf: a row of a helper field."""
        if "_flt" not in self.cpu.fields:
            self.create_field("_flt")
        r = self._flt_rows
        self._flt_rows += 1
        for c in cols:
            self.set_bit("_flt", r, c)
        if not cols:
            self._touch("_flt", 0)
        return abi.Row("_flt", r)

    # ---- mirror into the engine ----
    def sync(self):
        if not self.gpu:
            self._dirty = {k for k in self._dirty if k[0] == "cfg"}
            return
        for key in sorted(k for k in self._dirty if k[0] != "cfg"):
            field, view, shard = key
            fam_c = self.cpu.family(field, view)
            fam_g = self.gpu.family(field, view)
            data = self.orc.frag_serialize(fam_c, shard)
            if data:
                self.eng.frag_load(fam_g, shard, data)
                ids, counts = self.orc.rankings(fam_c, shard)
                if not self._cache.get(field, (False, 0))[0]:
                    self.eng.rank_set(fam_g, shard, ids, counts)
                else:
                    self.eng.rank_set(fam_g, shard, [], [])
        self._dirty = {k for k in self._dirty if k[0] == "cfg"}

    @property
    def ex(self) -> Executor:
        """The executor under test: the engine when backend == gpu, else the oracle itself."""
        self.sync()
        return self.gpu if self.gpu else self.cpu

    def close(self):
        if self.eng:
            self.eng.close()

    # ---- fragment-level calls, named like the reference's ----
    def frag_top(self, field: str, shard: int = 0, n: int = 0, src: Optional[abi.Call] = None, row_ids: Sequence[int] = (),
                 min_threshold: int = 0, tanimoto: int = 0) -> List[Tuple[int, int]]:
        ex = self.ex
        sp = ex.program(src) if src is not None else None
        return ex.be.topn_shard(ex.family(field), sp, n, shard, ids=list(row_ids), min_threshold=min_threshold, tanimoto=tanimoto)

    def frag_sum(self, field: str, filt: Optional[abi.Call] = None, shard: int = 0, bit_depth: Optional[int] = None):
        ex = self.ex
        bd = bit_depth if bit_depth is not None else ex.fields[field].bsig.bit_depth
        return ex.be.bsi_sum(ex.family(field, self._view(field)), bd, ex.program(filt) if filt is not None else None, [shard])

    def frag_minmax(self, field: str, is_max: bool, filt: Optional[abi.Call] = None, shard: int = 0, bit_depth: Optional[int] = None):
        ex = self.ex
        bd = bit_depth if bit_depth is not None else ex.fields[field].bsig.bit_depth
        v, c = ex.be.bsi_minmax_shards(ex.family(field, self._view(field)), bd, ex.program(filt) if filt is not None else None, [shard], is_max)
        return v[0], c[0]

    def frag_range(self, field: str, op: int, a: int, b: int = 0, shard: int = 0, bit_depth: Optional[int] = None) -> List[int]:
        """fragment.rangeOp / rangeBetween: raw predicate, no executor pre-checks."""
        ex = self.ex
        bd = bit_depth if bit_depth is not None else ex.fields[field].bsig.bit_depth
        call = abi.Call("Range", field_name=field, bsi_op=op, a=a, b=b, bit_depth=bd)
        return ex.be.eval_columns(ex.program(call), shard)

    def frag_rows(self, field: str, shard: int = 0, start: int = 0, column: Optional[int] = None, limit: Optional[int] = None) -> List[int]:
        """fragment.rows(start, filters...) of one shard."""
        ex = self.ex
        return ex.be.rows(ex.family(field), [shard], start=start, column=column, limit=limit)

    def import_positions(self, field: str, shard: int, set_pos: Sequence[int] = (), clear_pos: Sequence[int] = ()) -> None:
        """fragment.importPositions (the bulkImport path) on the oracle AND, independently, on the engine's resident copy."""
        self.sync()
        self.orc.frag_import_positions(self.cpu.family(field, self._view(field)), shard, list(set_pos), list(clear_pos))
        if self.gpu:
            self.eng.frag_apply(self.gpu.family(field, self._view(field)), shard, list(set_pos), list(clear_pos))

    def import_roaring(self, field: str, shard: int, data: bytes, clear: bool = False) -> int:
        """fragment.importRoaring on the oracle AND, independently, on the engine's resident copy."""
        self.sync()
        n = self.orc.frag_import_roaring(self.cpu.family(field, self._view(field)), shard, data, clear)
        if self.gpu:
            assert self.eng.frag_import_roaring(self.gpu.family(field, self._view(field)), shard, data, clear) == n
        return n
