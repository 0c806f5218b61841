"""Host-side mirror of the reference's executor for the hot path (executor.go).

``Executor`` keeps the reference's call surface — Count / Row / TopN / Sum / Min / Max over a
list of shards — and delegates every per-shard computation to a *backend* exposing the C-ABI
methods (``pilosa_b200.engine.Engine`` in production). It owns only what stays on the host
in the reference too: field metadata, BSI predicate re-basing (field.go:1526-1567,
executor.go:1536-1665) and the cross-device reduce (executor.go:2458-2611 mapReduce).

Multi-GPU: one process per GPU; shard ``s`` lives on rank ``s % world``;
each rank evaluates its own shards and only the tiny partials are exchanged with
``torch.distributed`` (NCCL over NVLink on the GPU box, gloo in the CPU tests):
Count/Sum -> all-reduce(sum); Min/Max -> all-gather of (val, count, position) triples and
the ValCount.smaller/larger rule; TopN -> all-gather of the pass-1 pairs, Pairs.Add, then
pass 2 with the merged candidate ids. Bitmap-returning calls need no exchange.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

from . import abi
from .abi import SHARD_WIDTH, Call, flatten

EXISTENCE_FIELD = "exists"  # index.go existenceFieldName


@dataclass
class BsiGroup:  # field.go:1505-1512
    min: int = 0
    max: int = 0
    base: int = 0
    bit_depth: int = 0

    def bit_depth_min(self) -> int:  # field.go:1579
        return self.base - (1 << self.bit_depth) + 1

    def bit_depth_max(self) -> int:  # field.go:1584
        return self.base + (1 << self.bit_depth) - 1

    def base_value(self, op: int, value: int) -> Tuple[int, bool]:  # field.go:1526-1552
        mn, mx = self.bit_depth_min(), self.bit_depth_max()
        bv = 0
        if op in (abi.BSI_GT, abi.BSI_GTE):
            if value > mx:
                return 0, True
            if value > mn:
                bv = value - self.base
        elif op in (abi.BSI_LT, abi.BSI_LTE):
            if value < mn:
                return 0, True
            bv = (mx - self.base) if value > mx else (value - self.base)
        elif op in (abi.BSI_EQ, abi.BSI_NEQ):
            if value < mn or value > mx:
                return 0, True
            bv = value - self.base
        return bv, False

    def base_value_between(self, lo: int, hi: int) -> Tuple[int, int, bool]:  # field.go:1555-1569
        mn, mx = self.bit_depth_min(), self.bit_depth_max()
        if hi < mn or lo > mx:
            return 0, 0, True
        return max(lo, mn) - self.base, min(hi, mx) - self.base, False


def bit_depth(v: int) -> int:  # field.go:1606-1613
    for i in range(63):
        if v < (1 << i):
            return i
    return 63


def bit_depth_int64(v: int) -> int:  # field.go:1616-1621
    return bit_depth(-v if v < 0 else v)


@dataclass
class Field:
    name: str
    type: str = "set"  # set | int | time
    families: Dict[str, int] = field(default_factory=dict)  # view -> family handle
    bsig: Optional[BsiGroup] = None
    time_quantum: str = ""  # "YMDH" subset for time fields (time.go:29-58)
    no_standard_view: bool = False  # FieldOptions.NoStandardView: only the time views are written


class Executor:
    def __init__(self, backend, rank: int = 0, world: int = 1, group=None):
        self.be = backend
        self.rank, self.world, self.group = rank, world, group
        self.fields: Dict[str, Field] = {}

    # ---- schema / loading (host bookkeeping; holder.go / field.go stay in Go) ----
    def create_field(self, name: str, type: str = "set", min: int = 0, max: int = 0, time_quantum: str = "",
                     no_standard_view: bool = False) -> Field:
        f = Field(name, type, time_quantum=time_quantum, no_standard_view=no_standard_view)
        if type == "int":
            # field.go:1047-1066: BitDepth from the declared range, Base 0 for new fields
            f.bsig = BsiGroup(min=min, max=max, base=0, bit_depth=builtins_max(bit_depth_int64(min), bit_depth_int64(max)))
            f.families["bsig_" + name] = self.be.family_create()
        elif not no_standard_view:
            f.families["standard"] = self.be.family_create()
        self.fields[name] = f
        return f

    def family(self, field_name: str, view: str = "standard") -> int:
        f = self.fields.get(field_name)
        if f is None:
            raise KeyError(f"field not found: {field_name}")  # ErrFieldNotFound
        if view not in f.families:
            f.families[view] = self.be.family_create()
        return f.families[view]

    def owns(self, shard: int) -> bool:
        return shard % self.world == self.rank

    def load_fragment(self, field_name: str, shard: int, data, view: Optional[str] = None) -> bool:
        """Loads the fragment on this rank if it owns the shard (shard % world == rank)."""
        if not self.owns(shard):
            return False
        f = self.fields[field_name]
        view = view or ("bsig_" + field_name if f.type == "int" else "standard")
        self.be.frag_load(self.family(field_name, view), shard, data)
        return True

    def local_shards(self, shards: Sequence[int]) -> List[int]:
        return [s for s in shards if self.owns(s)]

    def _local(self, shards: Sequence[int]):
        """local_shards(); a shard list passed as a TUPLE is marshalled to the C array once and cached
        (converting 8192 Python ints costs more than the query they address)."""
        if not isinstance(shards, tuple):
            return self.local_shards(shards)
        cache = self.__dict__.setdefault("_local_cache", {})
        hit = cache.get(shards)
        if hit is None:
            if len(cache) >= 32:
                cache.clear()
            lst = self.local_shards(shards)
            hit = cache[shards] = abi.u64_array(lst) if lst else lst
        return hit

    def _local_positions(self, shards: Sequence[int]):
        """Index of every local shard in the caller's list (cached for tuples, like _local)."""
        if not isinstance(shards, tuple):
            return [i for i, s in enumerate(shards) if self.owns(s)]
        cache = self.__dict__.setdefault("_position_cache", {})
        hit = cache.get(shards)
        if hit is None:
            if len(cache) >= 32:
                cache.clear()
            hit = cache[shards] = [i for i, s in enumerate(shards) if self.owns(s)]
        return hit

    # ---- call-tree helpers ----
    def row_time(self, field_name: str, row: int, from_time=None, to_time=None) -> Call:
        """Row(f=row, from=, to=) on a time field (executeRowShard, executor.go:1501-1531): the union of the row
        over the time-quantum views that tile [from, to) (viewsByTimeRange, time.go:104-181). Views the field does
        not have are skipped, as the reference skips a view without a fragment (:1519-1522); the result is Empty
        without a quantum or without any such view. The union is an ordinary n-ary Union of Row leaves, one per
        view family; beyond the 32 leaves of one device program the library evaluates it chunk-wise into temp rows."""
        from .timeviews import VIEW_STANDARD, views_by_time_range

        f = self.fields.get(field_name)
        if f is None:
            raise KeyError(f"field not found: {field_name}")
        if not f.time_quantum:
            return Call("Empty")
        if from_time is None:  # the zero time.Time (:1507-1508)
            from datetime import datetime
            from_time = datetime(1, 1, 1)
        if to_time is None:    # "current time + 1 day, in order to account for timezone differences" (:1509-1512)
            from datetime import datetime, timedelta
            to_time = datetime.now() + timedelta(days=1)
        rows = [Call("Row", field_name=field_name, row=row, view=v)
                for v in views_by_time_range(VIEW_STANDARD, from_time, to_time, f.time_quantum) if v in f.families]
        if not rows:
            return Call("Empty")
        return rows[0] if len(rows) == 1 else Call("Union", rows)

    def row_range(self, field_name: str, op: int, value: Optional[int] = None, value2: Optional[int] = None) -> Call:
        """Row(field <op> value): executeRowBSIGroupShard's host-side pre-checks
        (executor.go:1536-1665) reduced to a Range / NotNull / Empty leaf."""
        f = self.fields[field_name]
        g = f.bsig
        if g is None:
            raise KeyError("bsiGroup not found")  # ErrBSIGroupNotFound
        rng = lambda o, a=0, b=0: Call("Range", field_name=field_name, bsi_op=o, a=a, b=b, bit_depth=g.bit_depth)  # noqa: E731
        if op == abi.BSI_NEQ and value is None:  # != null
            return rng(abi.BSI_NOTNULL)
        if op == abi.BSI_BETWEEN:
            lo, hi, oor = g.base_value_between(value, value2)
            if oor:
                return abi.Empty()
            if value <= g.min and value2 >= g.max:
                return rng(abi.BSI_NOTNULL)
            return rng(abi.BSI_BETWEEN, lo, hi)
        bv, oor = g.base_value(op, value)
        if oor and op != abi.BSI_NEQ:
            return abi.Empty()
        if ((op == abi.BSI_LT and value > g.max) or (op == abi.BSI_LTE and value >= g.max) or
                (op == abi.BSI_GT and value < g.min) or (op == abi.BSI_GTE and value <= g.min)):
            return rng(abi.BSI_NOTNULL)
        if oor and op == abi.BSI_NEQ:
            return rng(abi.BSI_NOTNULL)
        return rng(op, bv)

    def program(self, call: Call):
        return flatten(call, lambda fld, view: self.family(fld, view))

    # ---- collectives ----
    def _allreduce_ints(self, vals: List[int]) -> List[int]:
        if self.world == 1:
            return vals
        import torch
        import torch.distributed as dist

        dev = "cuda" if dist.get_backend(self.group) == "nccl" else "cpu"
        t = torch.tensor(vals, dtype=torch.int64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return [int(x) for x in t.tolist()]

    def _allgather_ints(self, vals: List[int]) -> List[List[int]]:
        """The same-length int64 vector of every rank, by one tensor all_gather (no pickling: all_gather_object
        costs ~0.5 ms per call over NCCL, more than the kernels it merges)."""
        if self.world == 1:
            return [list(vals)]
        import torch
        import torch.distributed as dist

        dev = "cuda" if dist.get_backend(self.group) == "nccl" else "cpu"
        t = torch.tensor(vals, dtype=torch.int64, device=dev)
        outs = [torch.empty_like(t) for _ in range(self.world)]
        dist.all_gather(outs, t, group=self.group)
        return [[int(x) for x in o.tolist()] for o in outs]

    def _allgather_pairs(self, pairs: Sequence[Tuple[int, int]]) -> List[List[Tuple[int, int]]]:
        """(id, count) lists of every rank: lengths differ, so the longest is agreed on first and every rank
        sends [n, id0, count0, ...] padded to it (two small tensor collectives)."""
        if self.world == 1:
            return [list(pairs)]
        longest = max(r[0] for r in self._allgather_ints([len(pairs)]))
        if longest == 0:
            return [[] for _ in range(self.world)]
        flat = [len(pairs)]
        for i, c in pairs:
            flat += [_i64(i), _i64(c)]
        flat += [0] * (2 * (longest - len(pairs)))
        return [[(_u64(r[1 + 2 * k]), _u64(r[2 + 2 * k])) for k in range(r[0])] for r in self._allgather_ints(flat)]

    def _allgather(self, obj):
        if self.world == 1:
            return [obj]
        import torch.distributed as dist

        out = [None] * self.world
        dist.all_gather_object(out, obj, group=self.group)
        return out

    # ---- queries (executor.go) ----
    def count(self, call: Call, shards: Sequence[int]) -> int:  # executeCount :1793
        local = self._local(shards)
        n = self.be.count(self.program(call), local) if len(local) else 0
        return _u64(self._allreduce_ints([_i64(n)])[0])

    def count_batch(self, calls: Sequence[Call], shards: Sequence[int]) -> List[int]:
        local = self._local(shards)
        vals = [int(v) for v in self.be.count_batch([self.program(c) for c in calls], local)[: len(calls)]] if len(local) else [0] * len(calls)
        return [_u64(v) for v in self._allreduce_ints([_i64(v) for v in vals])]

    def row_columns(self, call: Call, shards: Sequence[int]) -> List[int]:
        """Row.Columns() of a bitmap call over shards (rank-local shards only; results stay sharded)."""
        prog = self.program(call)
        cols: List[int] = []
        for s in self.local_shards(sorted(shards)):
            cols += self.be.eval_columns(prog, s)
        return cols

    def row_segments(self, call: Call, shards: Sequence[int]) -> List[Tuple[int, bytes]]:
        """executeBitmapCall (executor.go:604-652) before the coordinator's Row.Merge: one
        (shard, Pilosa roaring bytes) segment per rank-local shard, in ascending shard order; the
        result stays sharded across ranks (no exchange for bitmap-returning calls)."""
        local = self.local_shards(sorted(shards))
        return list(zip(local, self.be.eval_roaring_shards(self.program(call), local)))

    def row_proto_columns(self, call: Call, shards: Sequence[int]) -> bytes:
        """The `Columns` field of internal.Row for the rank-local part of a bitmap call's result, as the reference's
        marshaller writes it (internal/public.pb.go:745-766 via encodeRow, encoding/proto/proto.go:408-457)."""
        return self.be.eval_proto_columns(self.program(call), self.local_shards(sorted(shards)))

    def sum(self, field_name: str, shards: Sequence[int], filter_call: Optional[Call] = None) -> Tuple[int, int]:  # executeSum :409
        f = self.fields[field_name]
        local = self._local(shards)
        fp = self.program(filter_call) if filter_call is not None else None
        s, c = self.be.bsi_sum(self.family(field_name, "bsig_" + field_name), f.bsig.bit_depth, fp, local) if len(local) else (0, 0)
        s, c = self._allreduce_ints([s, c])
        if c == 0:
            return 0, 0
        return _wrap_i64(s + c * f.bsig.base), c  # executor.go:718

    def _minmax(self, field_name: str, shards: Sequence[int], filter_call: Optional[Call], is_max: bool) -> Tuple[int, int]:
        f = self.fields[field_name]
        local = self._local(shards)
        fp = self.program(filter_call) if filter_call is not None else None
        if self.world == 1:  # one rank: the library folds ValCount.smaller / larger itself, in shard-list order
            if not len(local):
                return 0, 0
            fam = self.family(field_name, "bsig_" + field_name)
            v, c = (self.be.bsi_max if is_max else self.be.bsi_min)(fam, f.bsig.bit_depth, fp, local)
            return (v + f.bsig.base, c) if c else (0, 0)  # executor.go:759,797
        positions = self._local_positions(shards)  # places of the local shards in the caller's list
        best = None  # (val, count, position in the caller's shard list)
        if len(local):
            vals, counts = self.be.bsi_minmax_shards(self.family(field_name, "bsig_" + field_name), f.bsig.bit_depth, fp, local, is_max)
            for pos, v, c in zip(positions, vals, counts):
                if c == 0:
                    continue
                if best is None or (v > best[0] if is_max else v < best[0]):
                    best = (v, c, pos)
        mine = [1, best[0], _i64(best[1]), best[2]] if best is not None else [0, 0, 0, 0]
        cands = [(r[1], _u64(r[2]), r[3]) for r in self._allgather_ints(mine) if r[0]]
        if not cands:
            return 0, 0
        # ValCount.smaller/larger folded in shard-list order: the first shard attaining the extreme wins
        cands.sort(key=lambda b: ((-b[0] if is_max else b[0]), b[2]))
        v, c, _ = cands[0]
        return v + f.bsig.base, c  # executor.go:759,797

    def min(self, field_name, shards, filter_call=None):  # executeMin :445
        return self._minmax(field_name, shards, filter_call, False)

    def max(self, field_name, shards, filter_call=None):  # executeMax :481
        return self._minmax(field_name, shards, filter_call, True)

    def topn(self, field_name: str, shards: Sequence[int], n: int = 0, src: Optional[Call] = None, ids: Sequence[int] = (),
             threshold: int = 0, tanimoto: int = 0) -> List[Tuple[int, int]]:  # executeTopN :863
        fam = self.family(field_name)
        sp = self.program(src) if src is not None else None
        local = self._local(shards)
        if self.world == 1:
            return self.be.topn(fam, sp, n, local, ids=list(ids), min_threshold=threshold, tanimoto=tanimoto)
        p1 = self.be.topn(fam, sp, n, local, ids=list(ids), min_threshold=threshold, tanimoto=tanimoto, remote=True) if len(local) else []
        pairs = _merge_pairs(self._allgather_pairs(p1))
        if not pairs or ids:
            return pairs
        ids2 = sorted(i for i, _ in pairs)
        p2 = self.be.topn(fam, sp, n, local, ids=ids2, min_threshold=threshold, tanimoto=tanimoto, remote=True) if len(local) else []
        out = _merge_pairs(self._allgather_pairs(p2))
        return out[:n] if n else out


    def rows(self, field_name: str, shards: Sequence[int], previous: Optional[int] = None, limit: Optional[int] = None,
             column: Optional[int] = None, from_time=None, to_time=None) -> List[int]:  # executeRows :1277, executeRowsShard :1316-1441
        if column is not None:
            shards = [column // SHARD_WIDTH]  # :1291-1293
        local = self.local_shards(shards)
        start = 0 if previous is None else previous + 1
        f = self.fields[field_name]
        views = ["standard"]
        if f.type == "time" and (from_time is not None or to_time is not None or f.no_standard_view):  # :1340-1402
            from .timeviews import min_max_views, time_of_view, views_by_time_range

            if not f.time_quantum:
                return []
            mn, mx = min_max_views(list(f.families), f.time_quantum)
            if not mn or not mx:
                return []  # no time views yet
            min_t, max_t = time_of_view(mn, False), time_of_view(mx, True)
            from_time = min_t if from_time is None or from_time < min_t else from_time
            to_time = max_t if to_time is None or to_time > max_t else to_time
            views = views_by_time_range("standard", from_time, to_time, f.time_quantum)
        mine: List[int] = []
        for v in views:  # RowIDs.merge(viewRows, limit) over the views (:1430-1438): the smallest `limit` ids of the union
            if v not in f.families or not local:
                continue
            mine = sorted(set(mine) | set(self.be.rows(f.families[v], local, start=start, column=column, limit=limit)))
            if limit is not None:
                mine = mine[:limit]
        if self.world == 1:
            return mine
        merged = sorted(set().union(*[set(x or []) for x in self._allgather(mine)]))  # RowIDs.merge :1043
        return merged if limit is None else merged[:limit]

    def group_by(self, children: Sequence["RowsCall"], shards: Sequence[int], filter_call: Optional[Call] = None,
                 limit: Optional[int] = None, offset: Optional[int] = None) -> List[Tuple[Tuple[int, ...], int]]:
        """executeGroupBy (executor.go:1071-1158): [((row id per field), count)] in ascending group order."""
        if not children:
            raise ValueError("need at least one child call")  # :1073
        packed = []
        for ch in children:
            for name in (ch.field,):
                if name not in self.fields:
                    raise KeyError("field not found")  # ErrFieldNotFound :3104
            pre = None
            if ch.limit is not None or ch.column is not None:  # executed cluster-wide first (:1105-1116)
                pre = self.rows(ch.field, shards, previous=ch.previous, limit=ch.limit, column=ch.column)
                if not pre:
                    return []
            packed.append((self.family(ch.field), ch.previous, pre))
        fp = self.program(filter_call) if filter_call is not None else None
        local = self.local_shards(shards)
        mine = self.be.groupby(packed, fp, limit, local) if local else []
        if self.world > 1:  # mergeGroupCounts :1166 over the ranks' partial lists
            acc: Dict[Tuple[int, ...], int] = {}
            for lst in self._allgather(mine):
                for g, c in lst or []:
                    acc[g] = acc.get(g, 0) + c
            mine = sorted(acc.items())
            if limit is not None:
                mine = mine[:limit]
        if offset is not None and offset < len(mine):  # :1140-1146
            mine = mine[offset:]
        if limit is not None and limit < len(mine):    # :1148-1156
            mine = mine[:limit]
        return mine


@dataclass
class RowsCall:
    """A Rows() child of GroupBy / a Rows query: Rows(field[, previous=][, limit=][, column=])."""

    field: str
    previous: Optional[int] = None
    limit: Optional[int] = None
    column: Optional[int] = None


def builtins_max(a, b):
    return a if a > b else b


def _merge_pairs(lists) -> List[Tuple[int, int]]:  # Pairs.Add cache.go:356 + sort (count desc, id asc)
    m: Dict[int, int] = {}
    for lst in lists:
        for i, c in lst or []:
            m[i] = m.get(i, 0) + c
    return sorted(m.items(), key=lambda kv: (-kv[1], kv[0]))


def _i64(v: int) -> int:
    return v - (1 << 64) if v >= (1 << 63) else v


def _u64(v: int) -> int:
    return v & ((1 << 64) - 1)


def _wrap_i64(v: int) -> int:
    v &= (1 << 64) - 1
    return v - (1 << 64) if v >= (1 << 63) else v
