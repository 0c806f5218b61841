"""Bitmap-level known answers of roaring/roaring_test.go:462-913 and :1510 — Intersect / Union / UnionInPlace /
Difference / Xor over whole bitmaps, i.e. the key merge-joins of roaring.go:593-941 on top of the container kernels —
restated once as a table and run two ways:

 * on the oracle's Bitmap (`-m "not gpu"`): pins the CPU restatement against the reference's literals;
 * on the CUDA engine (`-m gpu`): a bitmap's values are fragment positions (row = v >> 20, column = v & (2^20 - 1)),
   so its serialized form IS a fragment file of shard 0. Each operand is loaded as its own field through
   pb200_frag_load and the bitmap-level op is the same op on every row, counts summed (pb200_count_batch) or
   columns concatenated (pb200_eval_columns).
"""
import random

import pytest

import oracle
from pilosa_b200.abi import Difference, Intersect, Row, Union, Xor

SW = 1 << 20

_cache = {}


def big():
    """628 .. 2683300, the 2.68 M-value operand several of the reference's tests build with a loop of Add calls."""
    if "big" not in _cache:
        _cache["big"] = list(range(628, 2683301))
    return _cache["big"]


def ref_bm_values():
    """testBM() (roaring_test.go:1326-1349): an array, a bitmap, a small run and a large run container; count 75007."""
    v = [(1 << 16) + i for i in range(0, 1024, 4)]
    v += [(2 << 16) + i for i in range(0, 16384, 2)]
    v += [(3 << 16) + i for i in range(1024)]
    v += [(4 << 16) + i for i in range(65535)]
    return v



def runs_25_8():  # TestBitmap_IntersectRunRun :711-722
    run_len, space = 25, 8
    offset = run_len // 2 + space
    return [offset + i + j for i in range(0, 65536 - run_len - offset, run_len + space) for j in range(run_len)]


def runs_32_1():  # :724-734
    run_len, space = 32, 1
    return [i + j for i in range(0, 65536 - run_len, run_len + space) for j in range(run_len)]


# (reference test, op, a, b, expected count, expected values or None)
CASES = [
    ("Intersection :462", "and", [0, 2683177], big, 1, None),
    ("Union1 :478 a", "or", [0, 2683177], lambda: big() + [4000000], 2682675, None),
    ("Union1 :478 b", "or", ref_bm_values, [0, 2683177], 75009, None),
    ("Union1 :478 c", "or", ref_bm_values, ref_bm_values, 75007, None),
    ("Intersection_Empty :639", "and", [0, 2683177], [], 0, []),
    ("IntersectArrayArray :650", "and", [0, 1, 7, 9, 11, 2683, 5005], [0, 2683, 2684, 5000], 2, [0, 2683]),
    ("IntersectArrayArray :650 reversed", "and", [0, 2683, 2684, 5000], [0, 1, 7, 9, 11, 2683, 5005], 2, [0, 2683]),
    ("IntersectBitmapBitmap :676", "and", list(range(0, 65536, 2)), list(range(0, 65536, 3)), 10923, None),
    ("IntersectRunRun :697 a", "and", [0, 1, 2, 3, 4, 5, 10, 11, 12, 13, 14, 15], [5, 6, 7, 8, 9, 10, 11], 3, [5, 10, 11]),
    ("IntersectRunRun :697 b", "and", runs_25_8, runs_32_1, 47628, None),
    ("Difference :738", "andnot", [0, 2683177], big, 1, [0]),
    ("Difference2 :750", "andnot", [0, 1, 2, 131072, 262144, SW + 5, SW + 7], [2, 3, 100000, 262144, 2 * SW + 1], 5,
     [0, 1, 131072, SW + 5, SW + 7]),
    ("Difference_Empty :759", "andnot", [0, 2683177], [], 2, [0, 2683177]),
    ("DifferenceArrayArray :768", "andnot", [0, 4, 8, 12, 16, 20], [1, 3, 6, 9, 12, 15, 18], 5, None),
    ("DifferenceArrayRun :777", "andnot", [0, 4, 8, 12, 16, 20, 36, 40, 44], [1, 2, 3, 4, 5, 6, 7, 8, 9, 30, 31, 32, 33, 34, 35, 36], 6, None),
    ("Union :788", "or", [0, 1000001, 1000002, 1000003], [0, 50000, 1000001, 1000002], 5, [0, 50000, 1000001, 1000002, 1000003]),
    ("Xor :821 a", "xor", [0, 1, 2, 3], ref_bm_values, 75011, None),
    ("Xor :821 b", "xor", ref_bm_values, [0, 1, 2, 3], 75011, None),
    ("Xor :821 c", "xor", ref_bm_values, ref_bm_values, 0, []),
    ("Xor_ArrayArray :840", "xor", [0, 1000001, 1000002, 1000003], [0, 50000, 1000001, 1000002], 2, [50000, 1000003]),
    ("Xor_Empty :857", "xor", [0, 50000, 1000001, 1000002], [], 4, None),
    ("Xor_ArrayBitmap :866 a", "xor", [1, 70, 200, 4097, 4098], list(range(0, 10000, 2)), 4999, None),
    ("Xor_ArrayBitmap :866 b", "xor", list(range(0, 10000, 2)), [1, 70, 200, 4097, 4098], 4999, None),
    ("Xor_ArrayBitmap :866 c", "xor", list(range(0, 10000, 2)), [], 5000, None),
    ("Xor_BitmapBitmap :896", "xor", list(range(1, 10000, 2)), list(range(0, 10000, 2)), 10000, None),
    ("Intersect :1510", "and", ref_bm_values, ref_bm_values, 75007, None),
]


def vals(x):
    return x() if callable(x) else x


# ---------------------------------------------------------------------------------------
# 1. the oracle's Bitmap against the reference's literals
# ---------------------------------------------------------------------------------------
def _oracle_op(op, a, b):
    return {"and": a.intersect, "or": a.union, "andnot": a.difference, "xor": a.xor}[op](b)


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_oracle_bitmap_level(case):
    name, op, a, b, count, expect = case
    bm_a, bm_b = oracle.Bitmap(*vals(a)), oracle.Bitmap(*vals(b))
    n_a, n_b = bm_a.count(), bm_b.count()
    res = _oracle_op(op, bm_a, bm_b)
    assert res.count() == count
    if expect is not None:
        assert res.slice() == expect
    assert (bm_a.count(), bm_b.count()) == (n_a, n_b)  # operands are not mutated (TestBitmap_UnionInPlace :812-819)
    if op == "and":
        assert bm_a.intersection_count(bm_b) == count   # Bitmap.IntersectionCount :568-590 is the same join
    if op == "xor":
        assert res.xor(res).count() == 0                # "equivalence" checks of :850-854, :884-888
    # the same operands after Optimize() (run containers where they pay off): the answer may not depend on the encoding
    bm_a.optimize(); bm_b.optimize()
    assert _oracle_op(op, bm_a, bm_b).count() == count


def test_oracle_union_in_place_n_way():
    """TestBitmap_UnionInPlace1 :505-549 / TestBitmap_UnionInPlace :798-819: the n-way union (unionInPlace roaring.go:737-886)
    into an empty target, operands unchanged."""
    bm0 = oracle.Bitmap(0, 2683177)
    bm1 = oracle.Bitmap(*(big() + [4000000]))
    assert oracle.Bitmap().union(bm0, bm1).count() == 2682675
    bm = oracle.Bitmap(*ref_bm_values())
    bm.optimize()
    assert oracle.Bitmap().union(bm, bm0).count() == 75009
    assert oracle.Bitmap().union(bm, bm).count() == 75007
    assert (bm0.count(), bm1.count()) == (2, 2682674)
    a, b = oracle.Bitmap(0, 1000001, 1000002, 1000003), oracle.Bitmap(0, 50000, 1000001, 1000002)
    assert oracle.Bitmap().union(a, b).count() == 5 and (a.count(), b.count()) == (4, 4)


def _prop_batches(rnd, n_batches):
    """The generator of TestBitmap_UnionInPlaceProp :551-637: batches of up to 100 random values below 10^6, now and
    then a run of 2 * 65536 consecutive values (a full container and its neighbours)."""
    out = []
    for _ in range(n_batches):
        s = set()
        if rnd.randrange(100) <= 2:
            start = rnd.randrange(1000000)
            s.update(range(start, start + 2 * 65536))
        s.update(rnd.randrange(1000000) for _ in range(rnd.randrange(100)))
        out.append(s)
    return out


def test_oracle_union_in_place_property():
    rnd = random.Random(20190416)
    for _ in range(60):
        sets = _prop_batches(rnd, rnd.randrange(100) + 2)
        bms = [oracle.Bitmap(*sorted(s)) for s in sets]
        got = bms[0].union(*bms[1:])
        assert got.slice() == sorted(set().union(*sets))


# ---------------------------------------------------------------------------------------
# 2. the CUDA engine: operands loaded as fragments of shard 0, the op applied per row
# ---------------------------------------------------------------------------------------
OPS = {"and": Intersect, "or": Union, "andnot": Difference, "xor": Xor}


class _EngineBitmaps:
    """One engine context; every distinct operand becomes a field whose shard-0 fragment is the serialized bitmap."""

    def __init__(self):
        from pilosa_b200.engine import Engine
        from pilosa_b200.executor import Executor

        self.eng = Engine(0)
        self.ex = Executor(self.eng)
        self.fields = {}

    def field(self, values):
        key = (len(values), hash(tuple(values[:64])), hash(tuple(values[-64:])), sum(values[::997]))
        if key not in self.fields:
            name = f"b{len(self.fields)}"
            self.ex.create_field(name)
            if values:  # an empty bitmap is a field without a fragment
                self.ex.load_fragment(name, 0, oracle.Bitmap(*values).serialize())
            self.fields[key] = (name, (max(values) >> 20) + 1 if values else 0)
        return self.fields[key]

    def op(self, op, operands):
        named = [self.field(v) for v in operands]
        n_rows = max(r for _, r in named)
        calls = [OPS[op](*[Row(name, r) for name, _ in named]) for r in range(n_rows)]
        return calls

    def count(self, op, operands):
        calls = self.op(op, operands)
        return sum(self.ex.count_batch(calls, [0])) if calls else 0

    def values(self, op, operands):
        out = []
        for r, call in enumerate(self.op(op, operands)):
            out += [(r << 20) + c for c in self.ex.row_columns(call, [0])]
        return out


@pytest.fixture(scope="module")
def gpu():
    g = _EngineBitmaps()
    yield g
    g.eng.close()


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_engine_bitmap_level(gpu, case):
    name, op, a, b, count, expect = case
    operands = [vals(a), vals(b)]
    assert gpu.count(op, operands) == count
    if expect is not None:
        assert gpu.values(op, operands) == expect
    elif count <= 100000:
        assert len(gpu.values(op, operands)) == count


@pytest.mark.gpu
def test_engine_n_way_union_property(gpu):
    """TestBitmap_UnionInPlaceProp :551-637 on the engine: the n-ary Union of 2 .. 101 operands, as the reference draws
    them (beyond 32 leaves the call is evaluated chunk-wise through temp rows), against the plain set union, values
    compared one by one."""
    rnd = random.Random(20190416)
    for _ in range(12):
        sets = [s for s in _prop_batches(rnd, rnd.randrange(100) + 2) if s]
        if len(sets) < 2:
            continue
        operands = [sorted(s) for s in sets]
        want = sorted(set().union(*sets))
        assert gpu.count("or", operands) == len(want)
        assert gpu.values("or", operands) == want
