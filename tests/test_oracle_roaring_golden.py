"""Pins the CPU oracle's container algebra against the reference's own golden vectors
(roaring/roaring_internal_test.go), extracted verbatim into tests/golden/*.json by
tests/golden/extract_reference_vectors.py. Set contents and cardinalities are compared."""
import json
import os

import pytest

import oracle
from oracle import ARRAY, BITMAP, RUN, Cont

G = os.path.join(os.path.dirname(__file__), "golden")
TABLES = json.load(open(os.path.join(G, "roaring_tables.json")))
COMBOS = json.load(open(os.path.join(G, "container_combinations.json")))["rows"]

W = 65536


def words_to_vals(words):
    if isinstance(words, dict):
        return sorted(bm(words).flat)
    return [i * 64 + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1]


def runs_to_pairs(runs):
    return [(r["start"], r["last"]) for r in (runs or [])]


def runs_to_vals(runs):
    out = []
    for s, l in runs_to_pairs(runs):
        out += list(range(s, l + 1))
    return out


def bm(x):
    """bitmap literal: list of words, or MakeBitmap(words) / named fixture call."""
    if isinstance(x, dict) and "call" in x:
        if x["call"] == "MakeBitmap":
            return Cont.bitmap_words(x["args"][0])
        name = x["call"]
        if name in FIXTURE_CALLS:
            return Cont.bitmap(PATTERNS[FIXTURE_CALLS[name]])
        name = name[len("bitmap"):]
        return Cont.bitmap(PATTERNS[name[0].lower() + name[1:]])
    return Cont.bitmap_words(x or [])


# ---- the ten named patterns of roaring_helpers_test.go:22-311 (restated) ----
PATTERNS = {
    "empty": [],
    "full": list(range(W)),
    "firstBitSet": [0],
    "lastBitSet": [W - 1],
    "firstBitUnset": list(range(1, W)),
    "lastBitUnset": list(range(W - 1)),
    "innerBitsSet": list(range(1, W - 1)),
    "outerBitsSet": [0, W - 1],
    "oddBitsSet": list(range(1, W, 2)),
    "evenBitsSet": list(range(0, W, 2)),
    "secondBitSet": [1],
    "lastBitFirstRowSet": [63],
    "firstBitSecondRowSet": [64],
}
FIXTURE_CALLS = {
    "bitmapFirstBitSet": "firstBitSet", "bitmapSecondBitSet": "secondBitSet",
    "bitmapLastBitFirstRowSet": "lastBitFirstRowSet", "bitmapFirstBitSecoundRowSet": "firstBitSecondRowSet",
    "bitmapFirstBitSecondRowSet": "firstBitSecondRowSet", "bitmapLastBitSet": "lastBitSet", "bitmapEmpty": "empty",
    "MakeLastBitSet": "lastBitSet",
}


def vals_to_runs(vals):
    runs = []
    for v in vals:
        if runs and runs[-1][1] + 1 == v:
            runs[-1][1] = v
        else:
            runs.append([v, v])
    return [tuple(r) for r in runs]


_CACHE = {}


def as_type(vals, typ):
    key = (id(vals), typ)
    if key not in _CACHE:
        _CACHE[key] = _as_type(vals, typ)
    return _CACHE[key]


def _as_type(vals, typ):
    if typ == ARRAY:
        return Cont.array(vals)
    if typ == BITMAP:
        return Cont.bitmap(vals)
    return Cont.run(vals_to_runs(vals))


OPS = {"intersect": oracle.OP_INTERSECT, "union": oracle.OP_UNION, "difference": oracle.OP_DIFFERENCE,
       "xor": oracle.OP_XOR, "flip": oracle.OP_FLIP}


@pytest.mark.parametrize("op", ["intersect", "union", "difference", "xor", "flip"])
def test_container_combinations(op):
    """roaring_internal_test.go:2713-3345 — 3x3 type pairs x 10x10 patterns, every op."""
    rows = [r for r in COMBOS if r[0] == op]
    assert rows
    n = 0
    for _, x, y, exp in rows:
        for tx in (ARRAY, BITMAP, RUN):
            for ty in ((ARRAY, BITMAP, RUN) if op != "flip" else (ARRAY,)):
                a = as_type(PATTERNS[x], tx)
                b = as_type(PATTERNS[y], ty) if op != "flip" else None
                typ, vals, _, _ = oracle.container_op(OPS[op], a, b)
                assert vals == PATTERNS[exp], f"{op}({tx}/{x}, {ty}/{y}) != {exp}"
                if op == "intersect":
                    assert oracle.container_icount(a, b) == len(PATTERNS[exp])
                n += 1
    assert n >= len(rows)


def test_intersection_count_known_answers():
    # roaring_internal_test.go:349-357
    a = Cont.array([1, 5, 10, 11, 12])
    b = Cont.run([(2, 10), (12, 13), (15, 16)])
    assert oracle.container_icount(a, b) == 3
    assert oracle.container_icount(b, a) == 3
    # :359-377
    words = [0] * 1024
    words[0] = 1 << 63
    assert oracle.container_icount(Cont.bitmap_words(words), Cont.run([(63, 64)])) == 1
    a = Cont.bitmap_words([0xF0000001, 0xFF00000000000000, 0xFF000000000000F0, 0x0F0000])
    b = Cont.run([(29, 31), (125, 134), (191, 197), (200, 300)])
    assert oracle.container_icount(a, b) == 14
    assert oracle.container_icount(b, a) == 14
    # :235-254 full x full in every representation
    full = PATTERNS["full"]
    for ta in (ARRAY, BITMAP, RUN):
        for tb in (ARRAY, BITMAP, RUN):
            assert oracle.container_icount(as_type(full, ta), as_type(full, tb)) == 65536


def test_intersection_count_tables():
    for t in TABLES["TestIntersectionCountRunRun"]:
        a, b = Cont.run(runs_to_pairs(t["aruns"])), Cont.run(runs_to_pairs(t["bruns"]))
        assert oracle.container_icount(a, b) == t["exp"]
    for t in TABLES["TestIntersectionCountArrayBitmap2"]:
        assert oracle.container_icount(Cont.array(t["array"]), Cont.bitmap_words(t["bitmap"])) == t["exp"]


def test_bitmap_count_range_and_runs():
    for t in TABLES["TestBitmapCountRange"]:
        assert oracle.container_stat(1, Cont.bitmap_words(t["bitmap"]), t["start"], t["end"]) == t["exp"]
    for t in TABLES["TestBitmapCountRuns"]:
        assert oracle.container_stat(0, Cont.bitmap_words(t["bitmap"])) == t["exp"]
    for t in TABLES["TestArrayCountRuns"]:
        if t["array"]:
            assert oracle.container_stat(0, Cont.array(t["array"])) == t["exp"]


def _check(op, a, b, exp_vals, exp_n=None):
    _, vals, _, _ = oracle.container_op(op, a, b)
    assert vals == sorted(exp_vals)
    if exp_n is not None:
        assert len(vals) == exp_n


def test_intersect_tables():
    for t in TABLES["TestIntersectArrayRun"]:
        _check(oracle.OP_INTERSECT, Cont.array(t["array"]), Cont.run(runs_to_pairs(t["runs"])), t["exp"] or [])
    for t in TABLES["TestIntersectRunRun"]:
        _check(oracle.OP_INTERSECT, Cont.run(runs_to_pairs(t["aruns"])), Cont.run(runs_to_pairs(t["bruns"])),
               runs_to_vals(t["exp"]), t["expN"])
    for t in TABLES["TestIntersectBitmapRunBitmap"]:
        _check(oracle.OP_INTERSECT, Cont.bitmap_words(t["bitmap"]), Cont.run(runs_to_pairs(t["runs"])),
               words_to_vals(t["exp"]), t["expN"])
    for t in TABLES["TestIntersectBitmapRunArray"]:
        _check(oracle.OP_INTERSECT, Cont.bitmap_words(t["bitmap"]), Cont.run(runs_to_pairs(t["runs"])), t["exp"], t["expN"])
    for t in TABLES["TestIntersectArrayBitmap"]:
        _check(oracle.OP_INTERSECT, Cont.array(t["array"]), bm(t["bitmap"]), t["exp"] or [])


def test_union_tables():
    for t in TABLES["TestUnionRunRun"]:
        _check(oracle.OP_UNION, Cont.run(runs_to_pairs(t["aruns"])), Cont.run(runs_to_pairs(t["bruns"])), runs_to_vals(t["exp"]))
    for t in TABLES["TestUnionArrayRun"]:
        _check(oracle.OP_UNION, Cont.array(t["array"]), Cont.run(runs_to_pairs(t["runs"])), t["exp"])
    for t in TABLES["TestUnionBitmapRun"]:
        _check(oracle.OP_UNION, Cont.bitmap_words(t["bitmap"]), Cont.run(runs_to_pairs(t["runs"])), words_to_vals(t["exp"]), t["expN"])


def test_difference_tables():
    for t in TABLES["TestDifferenceArrayRun"]:
        _check(oracle.OP_DIFFERENCE, Cont.array(t["array"]), Cont.run(runs_to_pairs(t["runs"])), t["exp"])
    for t in TABLES["TestDifferenceRunArray"]:
        _check(oracle.OP_DIFFERENCE, Cont.run(runs_to_pairs(t["runs"])), Cont.array(t["array"]), runs_to_vals(t["exp"]))
    for t in TABLES["TestDifferenceRunBitmap"]:
        _check(oracle.OP_DIFFERENCE, Cont.run(runs_to_pairs(t["runs"])), bm(t["bitmap"]), runs_to_vals(t["exp"]))
    for t in TABLES["TestDifferenceBitmapRun"]:
        _check(oracle.OP_DIFFERENCE, bm(t["bitmap"]), Cont.run(runs_to_pairs(t["runs"])), words_to_vals(t["exp"]))
    for t in TABLES["TestDifferenceBitmapArray"]:
        first_word = Cont.bitmap([v for v in bm(t["bitmap"]).flat if v < 64])  # the Go test slices bitmap[:1]
        _check(oracle.OP_DIFFERENCE, first_word, Cont.array(t["array"]), t["exp"])
    for t in TABLES["TestDifferenceBitmapBitmap"]:
        _check(oracle.OP_DIFFERENCE, Cont.bitmap_words(t["abitmap"]), Cont.bitmap_words(t["bbitmap"]), t["exp"])
    for t in TABLES["TestDifferenceRunRun"]:
        _check(oracle.OP_DIFFERENCE, Cont.run(runs_to_pairs(t["aruns"])), Cont.run(runs_to_pairs(t["bruns"])), runs_to_vals(t["exp"]))


def _lit(x):
    """NewContainerArray([...]) / NewContainerRun([...]) literals of TestXorArrayRun."""
    if x["call"] == "NewContainerArray":
        return Cont.array(x["args"][0] or [])
    if x["call"] == "NewContainerRun":
        return Cont.run(runs_to_pairs(x["args"][0]))
    raise KeyError(x["call"])


def test_xor_tables():
    for t in TABLES["TestXorArrayRun"]:
        exp = t["exp"]
        exp_vals = values_of(_lit(exp))
        _check(oracle.OP_XOR, _lit(t["a"]), _lit(t["b"]), exp_vals)
        _check(oracle.OP_XOR, _lit(t["b"]), _lit(t["a"]), exp_vals)
    for t in TABLES["TestXorRunRun"]:
        a, b = Cont.run(runs_to_pairs(t["aruns"])), Cont.run(runs_to_pairs(t["bruns"]))
        _check(oracle.OP_XOR, a, b, runs_to_vals(t["exp"]))
        _check(oracle.OP_XOR, b, a, runs_to_vals(t["exp"]))
    for t in TABLES["TestXorBitmapRun"]:
        _check(oracle.OP_XOR, Cont.bitmap_words(t["bitmap"]), Cont.run(runs_to_pairs(t["runs"])), words_to_vals(t["exp"]))


def values_of(c: Cont):
    _, vals, _, _ = oracle.container_op(oracle.OP_CONVERT, c, to_type=ARRAY)
    return vals


def test_conversion_tables():
    for t in TABLES["TestArrayToBitmap"]:
        typ, vals, _, _ = oracle.container_op(oracle.OP_CONVERT, Cont.array(t["array"]), to_type=BITMAP)
        assert vals == words_to_vals(t["exp"])
    for t in TABLES["TestBitmapToArray"]:
        typ, vals, _, _ = oracle.container_op(oracle.OP_CONVERT, Cont.bitmap_words(t["bitmap"]), to_type=ARRAY)
        assert vals == (t["exp"] or [])
    for t in TABLES["TestRunToBitmap"]:
        typ, vals, _, _ = oracle.container_op(oracle.OP_CONVERT, Cont.run(runs_to_pairs(t["runs"])), to_type=BITMAP)
        assert vals == words_to_vals(t["exp"])
    for t in TABLES["TestBitmapToRun"]:
        b = t["bitmap"]
        if isinstance(b, dict) and b["call"] == "make":  # tests[8]: zeros, then words 1022/1023 set (:1083-1084)
            b = [0] * 1022 + [0xFFFFFFFFFFFFFFFF] * 2
        elif isinstance(b, dict) and b["call"] == "getFullBitmap":
            b = [0xFFFFFFFFFFFFFFFF] * 1024
        typ, vals, rep, _ = oracle.container_op(oracle.OP_CONVERT, Cont.bitmap_words(b), to_type=RUN)
        assert vals == runs_to_vals(t["exp"])
        if vals:
            assert typ == RUN and rep == len(t["exp"])
    for t in TABLES["TestArrayToRun"]:
        typ, vals, rep, _ = oracle.container_op(oracle.OP_CONVERT, Cont.array(t["array"]), to_type=RUN)
        assert vals == runs_to_vals(t["exp"])
        if vals:
            assert rep == len(t["exp"])
    for t in TABLES["TestRunToArray"]:
        typ, vals, _, _ = oracle.container_op(oracle.OP_CONVERT, Cont.run(runs_to_pairs(t["runs"])), to_type=ARRAY)
        assert vals == (t["exp"] or [])


def test_shift_tables():
    for t in TABLES["TestShiftArray"]:
        _, vals, _, carry = oracle.container_op(oracle.OP_SHIFT, Cont.array(t["array"]))
        assert vals == (t["exp"] or [])
        if "carry" in t:
            assert carry == t["carry"]
    for t in TABLES["TestShiftBitmap"]:
        _, vals, _, carry = oracle.container_op(oracle.OP_SHIFT, bm(t["bitmap"]))
        exp = bm(t["exp"])
        assert vals == sorted(exp.flat)
        if "carry" in t:
            assert carry == t["carry"]
    for t in TABLES["TestShiftRun"]:
        _, vals, _, carry = oracle.container_op(oracle.OP_SHIFT, Cont.run(runs_to_pairs(t["runs"])))
        assert vals == runs_to_vals(t["exp"])
        # (the Go table's n/en fields are inconsistent with its own runs - {5..10} has 6 values, en says 5 -
        # and its failure condition ANDs them away, :3587; only runs and carry are pinned)
        assert carry == t["carry"]


def test_optimize_type_rule():
    # roaring.go:2289-2338
    typ, vals, rep, _ = oracle.container_op(oracle.OP_OPTIMIZE, Cont.array(list(range(100))))
    assert typ == RUN and rep == 1
    typ, *_ = oracle.container_op(oracle.OP_OPTIMIZE, Cont.bitmap(list(range(0, 8000, 2))))
    assert typ == ARRAY  # 4000 values, 4000 runs > n/2
    typ, *_ = oracle.container_op(oracle.OP_OPTIMIZE, Cont.array(list(range(0, 8192, 2))))
    assert typ == BITMAP  # n = 4096 is not < ArrayMaxSize
    typ, *_ = oracle.container_op(oracle.OP_OPTIMIZE, Cont.array([]))
    assert typ == oracle.NIL
    # run kept while runs <= n/2 and <= 2048
    typ, _, rep, _ = oracle.container_op(oracle.OP_OPTIMIZE, Cont.bitmap([v for s in range(0, 65536, 32) for v in (s, s + 1)]))
    assert typ == RUN and rep == 2048
