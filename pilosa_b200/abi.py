"""ctypes view of include/pilosa_b200.h: struct layouts, constants and the call-tree
flattener shared by the engine binding (pilosa_b200.engine) and the test oracle binding.

The call tree mirrors the reference's ``pql.Call`` (pql/ast.go:263-267) restricted to the
bitmap calls of the hot path; ``flatten`` produces the post-order ``pb200_node`` program
the C ABI takes (n-ary nodes are left folds, executor.go:1668-1733,1003-1025).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Callable, List, Optional, Sequence

SHARD_WIDTH = 1 << 20

OP_ROW, OP_INTERSECT, OP_UNION, OP_DIFFERENCE, OP_XOR, OP_NOT, OP_SHIFT, OP_BSI_RANGE, OP_EMPTY = range(1, 10)
BSI_EQ, BSI_NEQ, BSI_LT, BSI_LTE, BSI_GT, BSI_GTE, BSI_BETWEEN, BSI_NOTNULL = range(1, 9)
SUM_STRICT = 1

OK, ERR_INVALID, ERR_CUDA, ERR_FORMAT, ERR_NOMEM, ERR_NOTFOUND, ERR_RANGE, ERR_UNSUPPORTED = 0, -1, -2, -3, -4, -5, -6, -7


class Node(C.Structure):
    """pb200_node (include/pilosa_b200.h)."""

    _fields_ = [
        ("op", C.c_uint32),
        ("family", C.c_uint32),
        ("row", C.c_uint64),
        ("nargs", C.c_uint32),
        ("bsi_op", C.c_uint32),
        ("bit_depth", C.c_uint32),
        ("reserved", C.c_uint32),
        ("a", C.c_int64),
        ("b", C.c_int64),
    ]


class Pair(C.Structure):
    """pb200_pair — cache.go:337-341."""

    _fields_ = [("id", C.c_uint64), ("count", C.c_uint64)]


GROUPBY_MAX_FIELDS = 4
NO_LIMIT = (1 << 64) - 1
ROWS_COLUMN = 1


class GroupByChild(C.Structure):
    """pb200_groupby_child — one Rows() child of a GroupBy (executor.go:1091-1118)."""

    _fields_ = [("family", C.c_uint32), ("has_previous", C.c_uint32), ("previous", C.c_uint64),
                ("rows", C.POINTER(C.c_uint64)), ("n_rows", C.c_uint64)]


class Group(C.Structure):
    """pb200_group — GroupCount (executor.go:1160-1164) with row ids only."""

    _fields_ = [("rows", C.c_uint64 * GROUPBY_MAX_FIELDS), ("count", C.c_uint64)]


def pack_groupby_children(children):
    """children: [(family, previous or None, rows list or None)] -> (ctypes array, keep-alive list)."""
    arr = (GroupByChild * max(1, len(children)))()
    keep = []
    for i, (fam, prev, rows) in enumerate(children):
        arr[i].family = fam
        arr[i].has_previous = 0 if prev is None else 1
        arr[i].previous = 0 if prev is None else prev
        if rows:
            a = (C.c_uint64 * len(rows))(*rows)
            keep.append(a)
            arr[i].rows = C.cast(a, C.POINTER(C.c_uint64))
            arr[i].n_rows = len(rows)
    return arr, keep


def unpack_groups(out, n, k):
    return [(tuple(out[i].rows[j] for j in range(k)), out[i].count) for i in range(n)]


class Timing(C.Structure):
    _fields_ = [
        ("kernel_ms", C.c_double),
        ("h2d_ms", C.c_double),
        ("d2h_ms", C.c_double),
        ("algo_bytes", C.c_uint64),
        ("h2d_bytes", C.c_uint64),
        ("d2h_bytes", C.c_uint64),
        ("launches", C.c_uint32),
        ("reserved", C.c_uint32),
    ]


assert C.sizeof(Node) == 48 and C.sizeof(Pair) == 16 and C.sizeof(Timing) == 56


# ---------------------------------------------------------------------------------------
# Call tree (pql.Call restricted to bitmap calls)
# ---------------------------------------------------------------------------------------
@dataclass
class Call:
    name: str  # Row, Intersect, Union, Difference, Xor, Not, Shift, Range, Empty
    children: List["Call"] = field(default_factory=list)
    field_name: Optional[str] = None
    row: int = 0
    view: str = "standard"
    bsi_op: int = 0
    a: int = 0
    b: int = 0
    bit_depth: int = 0

    def __repr__(self) -> str:  # PQL-ish
        if self.name == "Row":
            return f"Row({self.field_name}={self.row})"
        if self.name == "Range":
            return f"Row({self.field_name} <op{self.bsi_op}> {self.a},{self.b})"
        return f"{self.name}({', '.join(map(repr, self.children))})"


def Row(field_name: str, row: int, view: str = "standard") -> Call:
    return Call("Row", field_name=field_name, row=row, view=view)


def Intersect(*children: Call) -> Call:
    return Call("Intersect", list(children))


def Union(*children: Call) -> Call:
    return Call("Union", list(children))


def Difference(*children: Call) -> Call:
    return Call("Difference", list(children))


def Xor(*children: Call) -> Call:
    return Call("Xor", list(children))


def Not(child: Call, existence_field: str = "exists") -> Call:
    return Call("Not", [child], field_name=existence_field)


def Shift(child: Call, n: int = 0) -> Call:
    """Shift(row, n=): Row.Shift (row.go:217-239) applies the one-step roaring shift n times; a missing
    n is 0 (executor.go:1772, IntArg of an absent key) and returns the row unchanged."""
    if n < 0:
        raise ValueError("cannot shift by negative values")  # row.go:219
    for _ in range(n):
        child = Call("Shift", [child])
    return child


def Empty() -> Call:
    return Call("Empty")


_OPS = {"Intersect": OP_INTERSECT, "Union": OP_UNION, "Difference": OP_DIFFERENCE, "Xor": OP_XOR}


def flatten(call: Call, family_of: Callable[[str, str], int]) -> "C.Array[Node]":
    """Post-order flattening. ``family_of(field, view)`` resolves a family handle."""
    out: List[Node] = []

    def rec(c: Call) -> None:
        if c.name == "Row":
            out.append(Node(op=OP_ROW, family=family_of(c.field_name, c.view), row=c.row))
        elif c.name == "Empty":
            out.append(Node(op=OP_EMPTY))
        elif c.name in _OPS:
            for ch in c.children:
                rec(ch)
            out.append(Node(op=_OPS[c.name], nargs=len(c.children)))
        elif c.name == "Not":
            rec(c.children[0])
            out.append(Node(op=OP_NOT, family=family_of(c.field_name, "standard"), row=0, nargs=1))
        elif c.name == "Shift":
            rec(c.children[0])
            out.append(Node(op=OP_SHIFT, nargs=1))
        elif c.name == "Range":
            out.append(
                Node(op=OP_BSI_RANGE, family=family_of(c.field_name, "bsig_" + c.field_name), bsi_op=c.bsi_op,
                     bit_depth=c.bit_depth, a=c.a, b=c.b)
            )
        else:
            raise ValueError(f"unknown call {c.name}")

    rec(call)
    return (Node * len(out))(*out)


def u64_array(vals: Sequence[int]) -> "C.Array[C.c_uint64]":
    if isinstance(vals, C.Array):  # prepared once by the caller (a 8192-entry list costs ~1 ms to convert)
        return vals
    return (C.c_uint64 * len(vals))(*vals)
