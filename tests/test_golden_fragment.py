"""Known-answer tests of the reference's fragment / executor tests, restated. Each test runs
against the CPU oracle (pins the oracle, `-m "not gpu"`) and against the CUDA engine through
the C ABI (`-m gpu`), with the expected values being the literals of the Go tests cited."""
import pytest

from pilosa_b200 import abi
from pilosa_b200.abi import Difference, Intersect, Not, Row, Shift, Union, Xor

from world import SW, World

BACKENDS = ["oracle", pytest.param("gpu", marks=pytest.mark.gpu)]


@pytest.fixture(params=BACKENDS)
def w(request):
    world = World(request.param)
    yield world
    world.close()


def _bsi(w, values, bit_depth=16, name="v"):
    w.create_field(name, type="int", min=-(1 << bit_depth) + 1, max=(1 << bit_depth) - 1)
    assert w.cpu.fields[name].bsig.bit_depth == bit_depth
    for col, v in values:
        w.set_value(name, col, v)


# ---- fragment_internal_test.go:373-423 TestFragment_Sum ----
def test_fragment_sum(w):
    _bsi(w, [(1000, 382), (2000, 300), (3000, 2818), (4000, 300)])
    assert w.frag_sum("v") == (3800, 4)
    assert w.frag_sum("v", w.new_row(2000, 4000, 5000)) == (600, 2)
    w.clear_value("v", 1000, 23)
    assert w.frag_sum("v") == (3800 - 382, 3)


# ---- fragment_internal_test.go:426-499 TestFragment_MinMax ----
def test_fragment_min_max(w):
    _bsi(w, [(1000, 382), (2000, 300), (3000, 2818), (4000, 300), (5000, 2818), (6000, 2817), (7000, 0)])
    mins = [(None, 0, 1), ((2000, 4000, 5000), 300, 2), ((2000, 4000), 300, 2), ((1,), 0, 0), ((1000,), 382, 1), ((7000,), 0, 1)]
    maxs = [(None, 2818, 2), ((2000, 4000, 5000), 2818, 1), ((2000, 4000), 300, 2), ((1,), 0, 0), ((1000,), 382, 1), ((7000,), 0, 1)]
    for cols, exp, cnt in mins:
        f = w.new_row(*cols) if cols else None
        assert w.frag_minmax("v", False, f) == (exp, cnt), ("min", cols)
    for cols, exp, cnt in maxs:
        f = w.new_row(*cols) if cols else None
        assert w.frag_minmax("v", True, f) == (exp, cnt), ("max", cols)


# ---- fragment_internal_test.go:502-694 TestFragment_Range ----
def test_fragment_range_eq_neq(w):
    _bsi(w, [(1000, 382), (2000, 300), (3000, 2818), (4000, 300)])
    assert w.frag_range("v", abi.BSI_EQ, 300) == [2000, 4000]
    assert w.frag_range("v", abi.BSI_NEQ, 300) == [1000, 3000]


def test_fragment_range_lt_gt_between(w):
    _bsi(w, [(1000, 382), (2000, 300), (3000, 2817), (4000, 301), (5000, 1), (6000, 0)])
    assert w.frag_range("v", abi.BSI_LT, 301) == [2000, 5000, 6000]
    assert w.frag_range("v", abi.BSI_LT, 300) == [5000, 6000]
    assert w.frag_range("v", abi.BSI_LTE, 301) == [2000, 4000, 5000, 6000]
    assert w.frag_range("v", abi.BSI_LTE, 300) == [2000, 5000, 6000]
    assert w.frag_range("v", abi.BSI_GT, 300) == [1000, 3000, 4000]
    assert w.frag_range("v", abi.BSI_GT, 301) == [1000, 3000]
    assert w.frag_range("v", abi.BSI_GTE, 300) == [1000, 2000, 3000, 4000]
    assert w.frag_range("v", abi.BSI_GTE, 301) == [1000, 3000, 4000]
    assert w.frag_range("v", abi.BSI_BETWEEN, 300, 2817) == [1000, 2000, 3000, 4000]
    assert w.frag_range("v", abi.BSI_BETWEEN, 301, 2817) == [1000, 3000, 4000]
    assert w.frag_range("v", abi.BSI_BETWEEN, 301, 2816) == [1000, 4000]
    assert w.frag_range("v", abi.BSI_BETWEEN, 300, 2816) == [1000, 2000, 4000]


# ---- fragment_internal_test.go:959-979 TestFragment_Top ----
def test_fragment_top(w):
    w.create_field("f")
    w.set_bits("f", 100, 1, 3, 200)
    w.set_bits("f", 101, 1)
    w.set_bits("f", 102, 1, 2)
    w.recalculate_cache("f")
    assert w.frag_top("f", n=2) == [(100, 3), (102, 2)]


# ---- fragment_internal_test.go:1017-1042 TestFragment_TopN_Intersect ----
def test_fragment_topn_intersect(w):
    w.create_field("f")
    src = w.new_row(1, 2, 3)
    w.set_bits("f", 100, 1, 10, 11, 12)
    w.set_bits("f", 101, 1, 2, 3, 4)
    w.set_bits("f", 102, 1, 2, 4, 5, 6)
    w.set_bits("f", 103, 1000, 1001, 1002)
    w.recalculate_cache("f")
    assert w.frag_top("f", n=3, src=src) == [(101, 3), (102, 2), (100, 1)]


# ---- fragment_internal_test.go:1044-1094 TestFragment_TopN_Intersect_Large ----
def test_fragment_topn_intersect_large(w):
    w.create_field("f")
    src = w.new_row(*range(980, 1000))
    fam = w._fam("f")
    w.orc.cache_config(fam, 0, False, 50000)
    for i in range(1000):
        for j in range(i):
            w.orc.set_bit(fam, 0, i, j)
    w.recalculate_cache("f")
    assert w.frag_top("f", n=10, src=src) == [(999 - k, 19 - k) for k in range(10)]


# ---- fragment_internal_test.go:1096-1113 TestFragment_TopN_IDs ----
def test_fragment_topn_ids(w):
    w.create_field("f")
    w.set_bits("f", 100, 1, 2, 3)
    w.set_bits("f", 101, 4, 5, 6, 7)
    w.set_bits("f", 102, 8, 9, 10, 11, 12)
    w._touch("f", 0)
    assert w.frag_top("f", row_ids=[100, 101, 200]) == [(101, 4), (100, 3)]


# ---- fragment_internal_test.go:1115-1131 TestFragment_TopN_NopCache (host never asks the engine:
# executor.go:978 rejects TopN on cache-less fields, so this one pins the oracle only) ----
def test_fragment_topn_nop_cache():
    w = World("oracle")
    w.create_field("f", cache_none=True)
    w.set_bits("f", 100, 1, 2, 3)
    w.set_bits("f", 101, 4, 5, 6, 7)
    assert w.frag_top("f", row_ids=[100, 101, 200]) == []


# ---- fragment_internal_test.go:1135-1195 TestFragment_TopN_CacheSize ----
def test_fragment_topn_cache_size(w):
    w.create_field("f", cache_size=3)
    w.set_bits("f", 100, 1, 2, 3)
    w.set_bits("f", 101, 4, 5, 6, 7)
    w.set_bits("f", 102, 8, 9, 10, 11, 12)
    w.set_bits("f", 103, 8, 9, 10, 11, 12, 13)
    w.set_bits("f", 104, 8, 9, 10, 11, 12, 13, 14)
    w.set_bits("f", 105, 10, 11)
    w.recalculate_cache("f")
    assert w.frag_top("f", n=5) == [(104, 7), (103, 6), (102, 5)]


# ---- executor_test.go:136-508: Difference / Intersect / Union / Xor / Count over two shards ----
def _two_shard_rows(w):
    w.create_field("general")
    w.set_bits("general", 10, 3, SW + 1)
    w.set_bits("general", 10, SW + 2)
    w.set_bits("general", 11, 2, SW + 2)
    return [0, 1]


def test_executor_difference_literal(w):
    # TestExecutor_Execute_Difference/RowIDColumnID (executor_test.go:136-152), literal
    w.create_field("general")
    w.set_bits("general", 10, 1, 2, 3)
    w.set_bits("general", 11, 2, 4)
    assert w.ex.row_columns(Difference(Row("general", 10), Row("general", 11)), [0]) == [1, 3]


def test_executor_set_ops(w):
    # same shapes as TestExecutor_Execute_Intersect/Union/Xor/Count (executor_test.go:214-508): two rows over two
    # shards; expected columns derived by hand from the set definitions
    shards = _two_shard_rows(w)
    ex = w.ex
    r10, r11 = Row("general", 10), Row("general", 11)
    assert ex.row_columns(Difference(r10, r11), shards) == [3, SW + 1]
    assert ex.row_columns(Intersect(r10, r11), shards) == [SW + 2]
    assert ex.row_columns(Union(r10, r11), shards) == [2, 3, SW + 1, SW + 2]
    assert ex.row_columns(Xor(r10, r11), shards) == [2, 3, SW + 1]
    assert ex.count(r10, shards) == 3
    assert ex.count(Union(), shards) == 0 and ex.row_columns(Xor(), shards) == []  # empty Union/Xor = empty row
    with pytest.raises(Exception, match="empty Intersect"):
        ex.count(Intersect(), shards)


# ---- executor_test.go:2759-2833 TestExecutor_Execute_Not ----
def test_executor_not(w):
    w.create_field("exists")
    w.create_field("f")
    for col in (1, 2, 3, SW + 1):
        w.set_bit("exists", 0, col)
    w.set_bits("f", 10, 3, SW + 1)
    w.set_bits("f", 11, 2)
    ex = w.ex
    assert ex.row_columns(Not(Row("f", 10)), [0, 1]) == [1, 2]
    assert ex.row_columns(Not(Row("f", 11)), [0, 1]) == [1, 3, SW + 1]
    assert ex.row_columns(Not(Union(Row("f", 10), Row("f", 11))), [0, 1]) == [1]
    assert ex.row_columns(Not(Row("f", 99)), [0, 1]) == [1, 2, 3, SW + 1]


# ---- executor_test.go:2713-2757 TestExecutor_Execute_Existence and :2759-2785 TestExecutor_Execute_Not/RowIDColumnID:
# the reference's own Set() calls (every set column also lands in the existence field) and expected columns ----
def test_executor_not_literal(w):
    w.create_field("exists")
    w.create_field("f")
    for col, row in ((3, 10), (SW + 1, 10), (SW + 2, 20)):
        w.set_bit("f", row, col)
        w.set_bit("exists", 0, col)
    ex = w.ex
    assert ex.row_columns(Row("f", 10), [0, 1]) == [3, SW + 1]
    assert ex.row_columns(Not(Row("f", 10)), [0, 1]) == [SW + 2]
    assert ex.row_columns(Not(Row("f", 20)), [0, 1]) == [3, SW + 1]
    assert ex.row_columns(Not(Row("f", 0)), [0, 1]) == [3, SW + 1, SW + 2]
    assert ex.row_columns(Not(Union(Row("f", 10), Row("f", 20))), [0, 1]) == []


# ---- executor_test.go:895-1215 TopN across two shards ----
def test_executor_topn_two_shards(w):
    w.create_field("f")
    w.create_field("other")
    # TestExecutor_Execute_TopN/RowIDColumnID (:895-935): literal Set() calls and expected pairs
    w.set_bits("f", 0, 0, 1, SW, SW + 2, 5 * SW + 100)
    w.set_bits("f", 10, 0, SW)
    w.set_bits("f", 20, SW)
    w.set_bits("other", 0, 0)
    for s in (0, 1, 5):
        w.recalculate_cache("f", s)
    assert w.ex.topn("f", [0, 1, 5], n=2) == [(0, 5), (10, 2)]


def test_executor_topn_src(w):
    # TestExecutor_Execute_TopN_Src (:1180-1215): literal bits and expected pairs
    w.create_field("f")
    w.create_field("other")
    w.set_bits("f", 0, 0, 1, SW)
    w.set_bits("f", 10, SW, SW + 1)
    w.set_bits("f", 20, SW, SW + 1, SW + 2)
    w.set_bits("other", 100, SW, SW + 1, SW + 2)
    for s in (0, 1):
        w.recalculate_cache("f", s)
    assert w.ex.topn("f", [0, 1], n=3, src=Row("other", 100)) == [(20, 3), (10, 2), (0, 1)]


# ---- executor_test.go:1122-1143 TestExecutor_Execute_TopN_fill: the second pass refetches the winner's count in the
# shard whose own top-1 was another row (executor.go:889-902) ----
def test_executor_topn_fill(w):
    w.create_field("f")
    w.set_bits("f", 0, 0, 1, 2, SW)
    w.set_bits("f", 1, SW + 2, SW)
    for s in (0, 1):
        w.recalculate_cache("f", s)
    assert w.ex.topn("f", [0, 1], n=1) == [(0, 4)]


# ---- executor_test.go:1146-1177 TestExecutor_Execute_TopN_fill_small: row 0 is nowhere a shard's top-1, yet wins ----
def test_executor_topn_fill_small(w):
    w.create_field("f")
    w.set_bits("f", 0, 0, SW, 2 * SW, 3 * SW, 4 * SW)
    w.set_bits("f", 1, 0, 1)
    w.set_bits("f", 2, SW, SW + 1)
    w.set_bits("f", 3, 2 * SW, 2 * SW + 1)
    w.set_bits("f", 4, 3 * SW, 3 * SW + 1)
    for s in range(5):
        w.recalculate_cache("f", s)
    assert w.ex.topn("f", [0, 1, 2, 3, 4], n=1) == [(0, 5)]


# ---- executor_test.go:1635-1685 TestExecutor_Execute_Row_Range/RowIDColumnID: timestamped Set() calls on a YMDH time
# field (each bit goes to the standard view and to the views of viewsByTime, time.go:91-101), Row(f=1, from=, to=) ----
def test_executor_row_time_range(w):
    from datetime import datetime, timedelta

    from pilosa_b200.timeviews import views_by_time

    for ex in w._execs():
        ex.create_field("f", type="time", time_quantum="YMDH")
    w._cache = getattr(w, "_cache", {})
    w._cache["f"] = (False, 50000)
    written = set()

    def set_ts(col, row, ts):
        for view in ["standard"] + views_by_time("standard", ts, "YMDH"):
            w.orc.set_bit(w.cpu.family("f", view), 0, row, col)
            w._dirty.add(("f", view, 0))
            written.add(view)

    ts = lambda s: datetime.strptime(s, "%Y-%m-%dT%H:%M")  # noqa: E731
    for col, when in ((2, "1999-12-31T00:00"), (3, "2000-01-01T00:00"), (4, "2000-01-02T00:00"), (5, "2000-02-01T00:00"),
                      (6, "2001-01-01T00:00"), (7, "2002-01-01T02:00")):
        set_ts(col, 1, ts(when))
    set_ts(8, 1, datetime.now() + timedelta(days=2))  # just outside the default end (now + 1 day)
    set_ts(2, 1, ts("1999-12-30T00:00"))
    set_ts(2, 1, ts("2002-02-01T00:00"))
    set_ts(2, 10, ts("2001-01-01T00:00"))

    def columns(frm, to):
        ex = w.ex
        return ex.row_columns(ex.row_time("f", 1, frm, to), [0])

    assert columns(ts("1999-12-31T00:00"), ts("2002-01-01T03:00")) == [2, 3, 4, 5, 6, 7]   # Standard
    assert columns(ts("1999-12-31T00:00"), None) == [2, 3, 4, 5, 6, 7]                      # From
    assert columns(None, ts("2002-01-01T02:00")) == [2, 3, 4, 5, 6]                         # To
    for view in sorted(written):  # Clear(2, f=1) clears the bit in the standard view and in the time views (Field.ClearBit, field.go:926-975)
        w.orc.clear_bit(w.cpu.family("f", view), 0, 1, 2)
        w._dirty.add(("f", view, 0))
    assert columns(ts("1999-12-31T00:00"), ts("2002-01-01T03:00")) == [3, 4, 5, 6, 7]      # Clear


# ---- executor_test.go:3286-3335 TestExecutor_Execute_RowsTime / RowsTimeEmpty: Rows(f, from=, to=) on a YMD time field
# without a standard view (executeRowsShard :1340-1402: the range is clamped to the existing views) ----
def test_executor_rows_time(w):
    from datetime import datetime

    from pilosa_b200.timeviews import views_by_time

    for ex in w._execs():
        ex.create_field("f", type="time", time_quantum="YMD", no_standard_view=True)
        ex.create_field("x", type="time", time_quantum="YMD", no_standard_view=True)
    w._cache = getattr(w, "_cache", {})
    w._cache["f"] = (False, 50000)
    ts = lambda s: datetime.strptime(s, "%Y-%m-%dT%H:%M")  # noqa: E731
    for col, row, when in ((9, 1, "2001-01-01T00:00"), (9, 2, "2002-01-01T00:00"), (9, 3, "2003-01-01T00:00"), (9, 4, "2004-01-01T00:00"),
                           (SW + 9, 13, "2003-02-02T00:00")):
        for view in views_by_time("standard", ts(when), "YMD"):
            w.orc.set_bit(w.cpu.family("f", view), col // SW, row, col)
            w._dirty.add(("f", view, col // SW))
    queries = [((ts("1999-12-31T00:00"), ts("2002-01-01T03:00")), [1]),
               ((ts("2002-01-01T00:00"), ts("2004-01-01T00:00")), [2, 3, 13]),
               ((ts("1990-01-01T00:00"), ts("1999-01-01T00:00")), []),
               ((None, None), [1, 2, 3, 4, 13]),
               ((ts("2002-01-01T00:00"), None), [2, 3, 4, 13]),
               ((None, ts("2003-02-03T00:00")), [1, 2, 3, 13]),
               ((ts("2002-01-01T00:00"), ts("2002-01-02T00:00")), [2])]
    for (frm, to), expected in queries:
        assert w.ex.rows("f", [0, 1], from_time=frm, to_time=to) == expected, (frm, to)
    assert w.ex.rows("x", [0, 1], from_time=ts("1999-12-31T00:00"), to_time=ts("2002-01-01T03:00")) == []  # RowsTimeEmpty


# ---- executor_test.go:1516-1633 Sum, :1262-1417 Min/Max (two shards, negatives) ----
def test_executor_sum_min_max(w):
    # TestExecutor_Execute_Sum/ColumnID (:1516-1565): literal Set() calls, Sum = 200/5, filtered 80/2
    w.create_field("x")
    w.create_field("foo", type="int", min=-990, max=1000)
    w.set_bits("x", 0, 0, SW + 1)
    for col, v in [(0, 20), (SW, 30), (SW + 2, 40), (5 * SW + 100, 50), (SW + 1, 60)]:
        w.set_value("foo", col, v)
    ex = w.ex
    shards = [0, 1, 5]
    assert ex.sum("foo", shards) == (200, 5)
    assert ex.sum("foo", shards, Row("x", 0)) == (80, 2)
    assert ex.sum("foo", shards, Row("x", 7)) == (0, 0)                  # executor.go:438-440
    assert ex.min("foo", [4]) == (0, 0)


def test_executor_min_max(w):
    # TestExecutor_Execute_MinMax/ColumnID (:1262-1325) + the Max table of the ColumnKey variant (:1385-1410)
    w.create_field("x")
    w.create_field("f", type="int", min=-1100, max=1000)
    w.set_bits("x", 0, 0, 3, SW + 1)
    w.set_bits("x", 1, 1)
    w.set_bits("x", 2, SW + 2)
    for col, v in [(0, 20), (1, -5), (2, -5), (3, 10), (SW, 30), (SW + 2, 40), (5 * SW + 100, 50), (SW + 1, 60)]:
        w.set_value("f", col, v)
    ex = w.ex
    shards = [0, 1, 5]
    for filt, exp, cnt in [(None, -5, 2), (Row("x", 0), 10, 1), (Row("x", 1), -5, 1), (Row("x", 2), 40, 1)]:
        assert ex.min("f", shards, filt) == (exp, cnt), ("min", repr(filt))
    for filt, exp, cnt in [(None, 60, 1), (Row("x", 0), 60, 1), (Row("x", 1), -5, 1), (Row("x", 2), 40, 1)]:
        assert ex.max("f", shards, filt) == (exp, cnt), ("max", repr(filt))


# ---- executor_test.go:1977-2168 TestExecutor_Execute_Row_BSIGroup (one shard + edges) ----
def test_executor_row_bsigroup(w):
    w.create_field("foo", type="int", min=-990, max=1000)
    w.create_field("edge", type="int", min=-900, max=1000)
    foo = [(50, 20), (SW, 20), (SW + 5, 30), (2, -10), (SW + 2, 60), (7, 0)]
    for col, v in foo:
        w.set_value("foo", col, v)
    for col, v in [(1, -100), (2, -900), (3, 1000), (4, 1)]:
        w.set_value("edge", col, v)
    ex = w.ex
    sh = [0, 1]

    def q(field, op, a=None, b=None):
        return ex.row_columns(ex.row_range(field, op, a, b), sh)

    by_val = lambda pred: sorted(c for c, v in foo if pred(v))  # noqa: E731
    assert q("foo", abi.BSI_EQ, 20) == by_val(lambda v: v == 20)
    assert q("foo", abi.BSI_NEQ, 20) == by_val(lambda v: v != 20)
    assert q("foo", abi.BSI_LT, 30) == by_val(lambda v: v < 30)
    assert q("foo", abi.BSI_LTE, 30) == by_val(lambda v: v <= 30)
    assert q("foo", abi.BSI_GT, 20) == by_val(lambda v: v > 20)
    assert q("foo", abi.BSI_GTE, 20) == by_val(lambda v: v >= 20)
    assert q("foo", abi.BSI_GT, -11) == by_val(lambda v: v > -11)
    assert q("foo", abi.BSI_LT, -5) == by_val(lambda v: v < -5)
    assert q("foo", abi.BSI_BETWEEN, 0, 30) == by_val(lambda v: 0 <= v <= 30)
    assert q("foo", abi.BSI_BETWEEN, -20, 20) == by_val(lambda v: -20 <= v <= 20)
    assert q("foo", abi.BSI_NEQ, None) == by_val(lambda v: True)                  # != null
    # out-of-range edges (executor.go:1640-1660)
    assert q("foo", abi.BSI_GT, 5000) == []
    assert q("foo", abi.BSI_LT, -5000) == []
    assert q("foo", abi.BSI_LT, 5000) == by_val(lambda v: True)                   # notNull
    assert q("foo", abi.BSI_GTE, -5000) == by_val(lambda v: True)
    assert q("foo", abi.BSI_EQ, 5000) == []
    assert q("foo", abi.BSI_NEQ, 5000) == by_val(lambda v: True)
    assert q("foo", abi.BSI_BETWEEN, 2000, 3000) == []
    assert q("foo", abi.BSI_BETWEEN, -5000, 5000) == by_val(lambda v: True)
    assert q("edge", abi.BSI_GTE, -900) == [1, 2, 3, 4]
    assert q("edge", abi.BSI_LTE, 1000) == [1, 2, 3, 4]
    assert q("edge", abi.BSI_EQ, -900) == [2] and q("edge", abi.BSI_EQ, 1000) == [3]


# ---- executor_test.go:3999-4085 TestExecutor_Execute_Shift ----
def test_executor_shift_bit0_and_container_boundary(w):
    w.create_field("general")
    w.set_bit("general", 10, 0)
    w.set_bit("general", 11, 65535)
    assert w.ex.row_columns(Shift(Row("general", 10), n=1), [0]) == [1]
    assert w.ex.row_columns(Shift(Shift(Row("general", 10), n=1), n=1), [0]) == [2]
    assert w.ex.row_columns(Shift(Row("general", 11), n=1), [0]) == [65536]


def test_executor_shift_shard_boundary(w):
    w.create_field("general")
    orig = [1, SW - 1, SW + 1]
    w.set_bits("general", 10, *orig)
    assert w.ex.row_columns(Shift(Row("general", 10), n=1), [0, 1]) == [2, SW, SW + 2]
    assert w.ex.row_columns(Shift(Row("general", 10), n=2), [0, 1]) == [3, SW + 1, SW + 3]
    assert w.ex.row_columns(Shift(Shift(Row("general", 10))), [0, 1]) == orig  # n defaults to 0
    assert w.ex.count(Shift(Row("general", 10), n=1), [0, 1]) == 3


def test_executor_shift_shard_boundary_no_create(w):
    w.create_field("general")
    w.set_bits("general", 10, SW - 2, SW - 1, SW, SW + 2)
    assert w.ex.row_columns(Shift(Row("general", 10), n=1), [0, 1]) == [SW - 1, SW, SW + 1, SW + 3]
    assert w.ex.row_columns(Shift(Shift(Row("general", 10), n=1), n=1), [0, 1]) == [SW, SW + 1, SW + 2, SW + 4]
    assert w.ex.count(Shift(Shift(Row("general", 10), n=1), n=1), [0, 1]) == 4


# The carried bit lives under the next shard's container key inside the SAME segment
# (roaring.go:970-974), so set ops see it only against other shifted rows: Intersect with a
# stored row drops it, Union / Xor / Difference keep it (row.go:106-213 work per segment).
def test_shift_overflow_in_set_ops(w):
    w.create_field("general")
    w.set_bits("general", 1, 5, SW - 1)
    w.set_bits("general", 2, 6, SW - 1, SW)
    w.set_bits("general", 3, 5, SW - 2, SW - 1)
    s1, s3 = Shift(Row("general", 1), n=1), Shift(Row("general", 3), n=1)
    for call, shards in [(Intersect(s1, Row("general", 2)), [0, 1]), (Union(s1, Row("general", 2)), [0, 1]), (Union(s1, Row("general", 2)), [0]),
                         (Xor(s1, s3), [0, 1]), (Difference(s3, s1), [0]), (Difference(Row("general", 2), s1), [0, 1]),
                         (Intersect(s1, s3), [0]), (Shift(Union(s1, s3), n=3), [0]), (Xor(Shift(s3, n=1), s1), [0])]:
        got = w.ex.row_columns(call, shards)
        assert got == w.cpu.row_columns(call, shards), repr(call)
        assert w.ex.count(call, shards) == len(got), repr(call)
    assert w.ex.row_columns(Intersect(s1, s3), [0]) == [6, SW]
    assert w.ex.row_columns(Union(s1, Row("general", 2)), [0, 1]) == [6, SW - 1, SW, SW]  # both segments hold column SW


# ---- executor_test.go:3244-3284 TestExecutor_Execute_Rows ----
def _import_bits(w, field, bits):
    if field not in w.cpu.fields:
        w.create_field(field)
    for row, col in bits:
        w.set_bit(field, row, col)


def test_executor_rows(w):
    _import_bits(w, "general", [(10, 0), (10, SW + 1), (11, 2), (11, SW + 2), (12, 2), (12, SW + 2), (13, 3)])
    sh = [0, 1]
    assert w.ex.rows("general", sh) == [10, 11, 12, 13]
    assert w.ex.rows("general", sh, limit=2) == [10, 11]
    assert w.ex.rows("general", sh, previous=10, limit=2) == [11, 12]
    assert w.ex.rows("general", sh, column=2) == [11, 12]
    assert w.ex.rows("general", sh, column=SW + 2, limit=1) == [11]
    assert w.ex.rows("general", sh, previous=13) == []
    assert w.ex.rows("general", [7]) == []


# ---- executor_test.go:3568-3890 TestExecutor_Execute_GroupBy ----
def test_executor_group_by(w):
    from pilosa_b200.executor import RowsCall as R

    sh = [0, 1, 2, 3]
    _import_bits(w, "general", [(10, 0), (10, 1), (10, SW + 1), (11, 2), (11, SW + 2), (12, 2), (12, SW + 2)])
    _import_bits(w, "sub", [(100, 0), (100, 1), (100, 3), (100, SW + 1), (110, 2), (110, 0)])
    with pytest.raises(ValueError, match="need at least one child call"):
        w.ex.group_by([], sh)
    with pytest.raises(KeyError):
        w.ex.group_by([R("missing")], sh)
    basic = [((10, 100), 3), ((10, 110), 1), ((11, 110), 1), ((12, 110), 1)]
    assert w.ex.group_by([R("general"), R("sub")], sh) == basic
    assert w.ex.group_by([R("general"), R("sub")], sh, filter_call=Row("general", 10)) == basic[:2]   # "Filter"
    assert w.ex.group_by([R("general", previous=10)], sh) == [((11,), 2), ((12,), 2)]                 # "check field offset no limit"
    assert w.ex.group_by([R("general", previous=10)], sh, limit=1) == [((11,), 2)]                    # "check field offset limit"
    assert w.ex.group_by([R("general")], sh, filter_call=Row("sub", 110)) == [((10,), 1), ((11,), 1), ((12,), 1)]
    # limit already trims the reduce (executor.go:1076-1081,1129), offset is applied to what is left (:1140-1146)
    assert w.ex.group_by([R("general"), R("sub")], sh, offset=1, limit=2) == basic[1:2]
    assert w.ex.group_by([R("general"), R("sub")], sh, offset=1) == basic[1:]

    _import_bits(w, "a", [(0, 1), (1, SW + 1)])
    _import_bits(w, "b", [(0, SW + 1), (1, 1)])
    assert w.ex.group_by([R("a"), R("b")], sh, limit=1) == [((0, 1), 1)]                              # "tricky data"

    same = [(0, 0), (0, 1), (0, 2), (1, 1), (2, 0), (2, 2), (3, 3)]
    for f in ("wa", "wb", "wc"):
        _import_bits(w, f, same)
    assert w.ex.group_by([R("wa"), R("wb"), R("wc", previous=1)], sh, limit=3) == \
        [((0, 0, 2), 2), ((0, 1, 0), 1), ((0, 1, 1), 1)]                                              # "test wrapping with previous"
    assert w.ex.group_by([R("wa", previous=3), R("wb", previous=3), R("wc", previous=3)], sh, limit=3) == []   # "previous is last result"
    assert w.ex.group_by([R("wa"), R("wb", previous=2), R("wc", previous=2)], sh, limit=1) == [((1, 0, 0), 1)]  # "test wrapping multiple"

    m = [(0, 0), (1, SW), (2, 0), (3, SW)]
    _import_bits(w, "ma", m)
    _import_bits(w, "mb", m)
    assert w.ex.group_by([R("ma"), R("mb")], sh, limit=5) == [((0, 0), 1), ((0, 2), 1), ((1, 1), 1), ((1, 3), 1), ((2, 0), 1)]
    assert w.ex.group_by([R("ma"), R("mb", limit=2)], sh, limit=5) == [((0, 0), 1), ((1, 1), 1), ((2, 0), 1), ((3, 1), 1)]
    assert w.ex.group_by([R("ma"), R("mb", column=SW)], sh, limit=5) == [((1, 1), 1), ((1, 3), 1), ((3, 1), 1), ((3, 3), 1)]

    n = [(0, 0), (0, SW), (1, 0), (1, SW)]
    _import_bits(w, "na", n)
    _import_bits(w, "nb", n)
    assert w.ex.group_by([R("na"), R("nb")], sh) == [((0, 0), 2), ((0, 1), 2), ((1, 0), 2), ((1, 1), 2)]   # "same rows in different shards"

    pp = [(0, 0), (1, 0), (2, 0), (3, 0), (3, 91000), (3, SW), (3, SW * 2), (3, SW * 3)]
    for f in ("ppa", "ppb", "ppc"):
        _import_bits(w, f, pp)
    total = w.ex.group_by([R("ppa"), R("ppb"), R("ppc")], sh, limit=3)                                # paging with previous
    res = total
    while len(total) < 64:
        last = res[-1][0]
        res = w.ex.group_by([R("ppa", previous=last[0]), R("ppb", previous=last[1]), R("ppc", previous=last[2])], sh, limit=3)
        assert res, "paging stalled"
        total += res
    exp = [((i // 16, (i % 16) // 4, i % 4), 1) for i in range(64)]
    exp[63] = ((3, 3, 3), 5)
    assert total == exp


# ---- fragment_internal_test.go:1463-1509 TestFragment_Tanimoto / TestFragment_Zero_Tanimoto ----
def test_fragment_tanimoto(w):
    w.create_field("f")
    src = w.new_row(1, 2, 3)
    w.set_bits("f", 100, 1, 3, 2, 200)
    w.set_bits("f", 101, 1, 3)
    w.set_bits("f", 102, 1, 2, 10, 12)
    w.recalculate_cache("f")
    assert w.frag_top("f", src=src, tanimoto=50) == [(100, 3), (101, 2)]
    assert w.frag_top("f", src=src, tanimoto=0) == [(100, 3), (101, 2), (102, 2)]


# ---- fragment_internal_test.go:2634-2712 TestFragment_RowsIteration ----
def test_fragment_rows_iteration(w):
    w.create_field("f")
    for i in range(100, 200):
        w.set_bit("f", i, i % 2)
    assert w.frag_rows("f") == list(range(100, 200))
    assert w.frag_rows("f", column=1) == list(range(101, 200, 2))
    w.create_field("g")
    w.set_bit("g", 1, 66000)
    w.set_bit("g", 2, 66000)
    w.set_bit("g", 2, 166000)
    assert w.frag_rows("g") == [1, 2] and w.frag_rows("g", column=66000) == [1, 2]
    w.create_field("h")   # "combinations", thinned out: every new bit keeps both listings complete
    expected = []
    for r in range(1, 10000, 1100):
        expected.append(r)
        for c in range(1, SW - 1, 210000):
            w.set_bit("h", r, c)
            assert w.frag_rows("h") == expected
            assert w.frag_rows("h", column=c) == expected


# ---- fragment_internal_test.go:2715-2760 TestFragment_RoaringImport ----
def test_fragment_roaring_import(w):
    import oracle

    tests = [[[0], [1]],
             [[0, 65535, 65536, 65537, 65538, 65539, 130000], [1000, 67000, 130000]],
             [[0, 65535, 65536, 65537, 65538, 65539, 130000], [0, 65535, 65536, 65537, 65538, 65539, 130000]],
             [[0, 65535, 65536, (1 << 20) + 1, (1 << 20) * 2 + 1], [1, (1 << 20) + 2, (1 << 20) * 2 + 2]]]
    for i, test in enumerate(tests):
        name = f"imp{i}"
        w.create_field(name)
        w.set_bit(name, 9, 9)  # the fragment exists on both sides before the first import
        seen = set()
        for positions in test:
            changed = w.import_roaring(name, 0, oracle.Bitmap(*positions).serialize(), False)
            assert changed == len(set(positions) - seen)
            seen |= set(positions)
            by_row = {}
            for p in sorted(seen):  # calcExpected: position -> (row, column)
                by_row.setdefault(p >> 20, []).append(p & (SW - 1))
            for row, cols in by_row.items():
                assert w.ex.row_columns(Row(name, row), [0]) == cols, (i, row)
        removed = w.import_roaring(name, 0, oracle.Bitmap(*test[0]).serialize(), True)  # and clearing them again
        assert removed == len(set(test[0]))
        assert sum(w.ex.count(Row(name, r), [0]) for r in range(3)) == len(seen - set(test[0]))


# ---- fragment.importRoaring's cache update (fragment.go:2258-2276) — no reference test looks at the cache after an
# import, so the expectations below are worked out by hand from that code and rankCache (cache.go:184-192 BulkAdd,
# :245-281 recalculate) for a cache of 4 entries (thresholdBuffer = int(1.1 * 4) = 4). Pins the oracle; the engine's
# own rank-cache state is compared with the oracle step by step in test_gpu_parity.py::test_import_roaring_rank_recount.
def test_import_roaring_cache_update_by_hand():
    import oracle

    def frag(rows):
        return oracle.Bitmap(*sorted(r * SW + c for r, cols in rows.items() for c in cols)).serialize()

    orc = oracle.Oracle(workers=1)
    fam = orc.family_create()
    orc.cache_config(fam, 0, False, 4)
    # rows 0..7 with 10, 17, .. 59 bits; openCache + Recalculate: top 4 ranked, threshold = count of the 5th (row 3: 31),
    # 8 entries > 4 -> the unranked ones leave the map
    orc.frag_load(fam, 0, frag({r: range(10 + 7 * r) for r in range(8)}))
    assert orc.rankings(fam, 0) == ([7, 6, 5, 4], [59, 52, 45, 38])
    steps = [
        # cached row 5: 45 + 20
        ({5: range(1000, 1020)}, False, 20, ([5, 7, 6, 4], [65, 59, 52, 38])),
        # row 1 is not cached: Get() = 0, 0 + 5 < 31 -> BulkAdd drops it
        ({1: range(1000, 1005)}, False, 5, ([5, 7, 6, 4], [65, 59, 52, 38])),
        # row 2 is not cached either: 0 + 60 >= 31 -> enters with 60 (it holds 84 bits); 5 entries: threshold = 38, row 4 leaves
        ({2: range(1000, 1060)}, False, 60, ([5, 2, 7, 6], [65, 60, 59, 52])),
        # a new row of 200 bits: threshold = 52, row 6 leaves the map
        ({30: range(200)}, False, 200, ([30, 5, 2, 7], [200, 65, 60, 59])),
        # ... cleared again: 200 - 200 = 0 < 52 -> dropped by BulkAdd, the cached 200 stays
        ({30: range(200)}, True, 200, ([30, 5, 2, 7], [200, 65, 60, 59])),
        # clear on uncached row 0: 0 + uint64(-3) wraps and passes the threshold; threshold = 59, row 7 leaves
        ({0: range(3)}, True, 3, ([0, 30, 5, 2], [2**64 - 3, 200, 65, 60])),
        # bits already set: no row changed, no Recalculate
        ({5: range(5)}, False, 0, ([0, 30, 5, 2], [2**64 - 3, 200, 65, 60])),
        # row 6 left the map two steps ago: 0 - 20 wraps as well
        ({6: range(20)}, True, 20, ([0, 6, 30, 5], [2**64 - 3, 2**64 - 20, 200, 65])),
    ]
    for rows, clear, changed, ranking in steps:
        assert orc.frag_import_roaring(fam, 0, frag(rows), clear) == changed
        assert orc.rankings(fam, 0) == ranking
    # importPositions (fragment.go:2051-2098, the bulkImport path): the rows named by the positions get their count
    # re-read from storage. The map holds {0, 6, 30, 5} and the threshold is 60 (row 2's cached count when it left).
    W3, W20 = 2**64 - 3, 2**64 - 20
    pos_steps = [
        # row 2 really holds 84 bits: one more = 85 >= 60 -> back in; threshold = 65, row 5 leaves
        ([2 * SW + 5000], [], 1, ([0, 6, 30, 2], [W3, W20, 200, 85])),
        # the last 7 bits of row 0 cleared: count 0 < 65 -> dropped, the wrapped count stays; 4 entries: threshold = 1
        ([], [c for c in range(3, 10)], 7, ([0, 6, 30, 2], [W3, W20, 200, 85])),
        # rows 1 and 6 named: both re-read (23 and 31 bits), row 6 loses its wrapped count; threshold = 23, row 1 leaves
        ([1 * SW + 9000], [6 * SW + 20], 2, ([0, 30, 2, 6], [W3, 200, 85, 31])),
        # a position that is already set still names its row: row 7 (59 bits) is re-read and enters
        ([7 * SW + 0], [], 0, ([0, 30, 2, 7], [W3, 200, 85, 59])),
    ]
    for set_pos, clear_pos, changed, ranking in pos_steps:
        assert orc.frag_import_positions(fam, 0, set_pos, clear_pos) == changed
        assert orc.rankings(fam, 0) == ranking


# ---- cache_test.go:24-35 TestCache_Rank: 5 ids with the same count into a cache of 3; Recalculate cuts the entries ----
def test_cache_rank():
    import oracle

    orc = oracle.Oracle(workers=1)
    fam = orc.family_create()
    cache_size = 3
    orc.cache_config(fam, 0, False, cache_size)
    for i in range(1, 2 * cache_size):
        orc.cache_add(fam, 0, i, 3)
    orc.recalculate_cache(fam, 0)
    assert orc.cache_len(fam, 0) == cache_size
    assert orc.rankings(fam, 0)[1] == [3, 3, 3]


# ---- fragment_internal_test.go:2761-2833 TestFragment_RoaringImportTopN: bulkImport twice, then importRoaring, f.top()
# after each step must rank every row by its new count — the rank-cache updates of :2080-2093 and :2258-2276 ----
def test_fragment_roaring_import_topn(w):
    import oracle

    w.create_field("f")
    pos = lambda rows, cols: [r * SW + c for r, c in zip(rows, cols)]  # noqa: E731
    w.import_positions("f", 0, pos([4, 4, 4, 4], [0, 1, 2, 3]))
    assert w.frag_top("f") == [(4, 4)]
    w.import_positions("f", 0, pos([5, 5, 5, 5, 5], [0, 1, 2, 3, 4]))
    assert w.frag_top("f") == [(5, 5), (4, 4)]
    w.import_roaring("f", 0, oracle.Bitmap(0, 65535, 65536, (1 << 20) + 1, (1 << 20) + 2, (1 << 20) * 2 + 1).serialize())
    assert w.frag_top("f") == [(5, 5), (4, 4), (0, 3), (1, 2), (2, 1)]


# ---- fragment_internal_test.go:1570-1685 TestFragment_ImportSet: bulkImport of (row, column) pairs, then a bulkImport
# with Clear; the table literally (importPositions = pb200_frag_apply on the engine) ----
def test_fragment_import_set(w):
    table = [
        ([1, 1, 1, 1], [0, 1, 2, 3], {1: [0, 1, 2, 3]}, [], [], {1: [0, 1, 2, 3]}),
        ([1, 1, 1, 1, 2, 2, 2, 2], [0, 1, 2, 3, 0, 1, 2, 3], {1: [0, 1, 2, 3], 2: [0, 1, 2, 3]}, [1, 1, 2], [1, 2, 3], {1: [0, 3], 2: [0, 1, 2]}),
        ([1, 1, 1, 1, 2], [0, 1, 2, 3, 1], {1: [0, 1, 2, 3], 2: [1]}, [1, 1, 1, 1], [0, 1, 2, 3], {1: [], 2: [1]}),
        ([1, 1, 1, 1, 2, 2, 1], [0, 1, 2, 3, 1, 8, 1], {1: [0, 1, 2, 3], 2: [1, 8]}, [1, 1], [0, 0], {1: [1, 2, 3], 2: [1, 8]}),
        ([1, 2, 3], [8, 8, 8], {1: [8], 2: [8], 3: [8]}, [1, 2, 3], [9, 9, 9], {1: [8], 2: [8], 3: [8]}),
    ]
    for i, (set_rows, set_cols, set_exp, clr_rows, clr_cols, clr_exp) in enumerate(table):
        name = f"is{i}"
        w.create_field(name)
        w.import_positions(name, 0, [r * SW + c for r, c in zip(set_rows, set_cols)])
        for row, cols in set_exp.items():
            assert w.ex.row_columns(Row(name, row), [0]) == cols, (i, "set", row)
        w.import_positions(name, 0, [], [r * SW + c for r, c in zip(clr_rows, clr_cols)])
        for row, cols in clr_exp.items():
            assert w.ex.row_columns(Row(name, row), [0]) == cols, (i, "clear", row)


# ---- fragment_internal_test.go:3099-3178 TestFragmentPositionsForValue: the BSI bit layout (exists row 0, sign row 1,
# magnitude bit i in row 2 + i; fragment.go:934-972). Oracle only: a value written with setValue must leave exactly the
# reference's toSet positions set and its toClear positions clear — the layout every BSI kernel reads. ----
def test_fragment_positions_for_value():
    import oracle

    table = [  # (column, bit depth, value, toSet, toClear)
        (0, 1, 0, [0], [SW, SW * 2]),
        (0, 3, 0, [0], [SW * 1, SW * 2, SW * 3, SW * 4]),
        (1, 3, 0, [1], [SW + 1, SW * 2 + 1, SW * 3 + 1, SW * 4 + 1]),
        (0, 1, 1, [0, SW * 2], [SW]),
        (0, 4, 10, [0, SW * 3, SW * 5], [SW * 1, SW * 2, SW * 4]),
        (0, 5, 10, [0, SW * 3, SW * 5], [SW * 1, SW * 2, SW * 4, SW * 6]),
    ]
    for i, (col, depth, value, to_set, to_clear) in enumerate(table):
        orc = oracle.Oracle(workers=1)
        fam = orc.family_create()
        for p in to_clear:  # pre-set what must be cleared, so that "clear" is observable
            orc.set_bit(fam, 0, p // SW, p % SW)
        orc.set_value(fam, 0, col, depth, value)
        have = set(oracle.Bitmap.deserialize(orc.frag_serialize(fam, 0)).slice())
        assert have == set(to_set), (i, sorted(have))
    # a negative value sets the sign row as well (:953-957)
    orc = oracle.Oracle(workers=1)
    fam = orc.family_create()
    orc.set_value(fam, 0, 7, 4, -10)
    assert set(oracle.Bitmap.deserialize(orc.frag_serialize(fam, 0)).slice()) == {7, SW + 7, SW * 3 + 7, SW * 5 + 7}


# ---- fragment_internal_test.go:222-368 TestFragment_SetValue (OK / Overwrite / Clear / NotExists / QuickCheck): the BSI
# write path of the oracle, which every Sum / Min / Max / Range fixture of this suite is built with. Oracle only. ----
def test_fragment_set_value():
    import random

    import oracle

    def frag():
        orc = oracle.Oracle(workers=1)
        return orc, orc.family_create()

    orc, fam = frag()                                   # OK
    assert orc.set_value(fam, 0, 100, 16, 3829) is True
    assert orc.value(fam, 0, 100, 16) == (3829, True)
    assert orc.set_value(fam, 0, 100, 16, 3829) is False
    orc, fam = frag()                                   # Overwrite
    assert orc.set_value(fam, 0, 100, 16, 3829) is True
    assert orc.set_value(fam, 0, 100, 16, 2028) is True
    assert orc.value(fam, 0, 100, 16) == (2028, True)
    orc, fam = frag()                                   # Clear
    assert orc.set_value(fam, 0, 100, 16, 3829) is True
    assert orc.clear_value(fam, 0, 100, 16, 2028) is True
    assert orc.value(fam, 0, 100, 16) == (0, False)
    orc, fam = frag()                                   # NotExists
    assert orc.set_value(fam, 0, 100, 10, 20) is True
    assert orc.value(fam, 0, 101, 11) == (0, False)
    rnd = random.Random(222)                            # QuickCheck, with the reference's bounds
    for _ in range(40):
        depth, bit_n = rnd.randrange(62) + 1, rnd.randrange(99) + 1
        orc, fam = frag()
        want = {}
        for _ in range(rnd.randrange(50)):
            v = rnd.getrandbits(64) % (1 << depth)
            want[v % bit_n] = v
            orc.set_value(fam, 0, v % bit_n, depth, v)
        for col, v in want.items():
            assert orc.value(fam, 0, col, depth) == (v, True), (depth, col)
