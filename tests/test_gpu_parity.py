"""Parity tests proper: the CUDA engine, called through the C ABI, against the CPU oracle on
the same seeded inputs (bit-exact: all integer work). Run on the B200 box with -m gpu."""
import random

import pytest

import oracle
from pilosa_b200 import abi, synth
from pilosa_b200.abi import Difference, Empty, Intersect, Not, Row, Shift, Union, Xor
from pilosa_b200.executor import Executor

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", params=["single", "multi", "multi-fanout"])
def pair(request):
    """Engine + oracle executors over the same small synthetic index (C1/C2/C3/C4/C5 shapes). The whole module runs
    three times: on one context, and through the pb200_multi handle over every visible GPU — once with its router
    (these small calls then run on ONE device over the peers' HBM) and once forced onto the shards' owners."""
    from pilosa_b200.engine import Engine, MultiEngine, device_count

    if request.param == "single":
        eng = Engine(0)
    else:
        eng = MultiEngine(list(range(device_count())))
        if request.param == "multi-fanout":
            eng.set_option(MultiEngine.OPT_SMALL_CALL_BYTES, 0)
    orc = oracle.Oracle(workers=8)
    g, c = Executor(eng), Executor(orc)
    shards = [0, 1, 2, 5]  # shard 3/4 absent on purpose (missing fragments)
    for ex in (g, c):
        ex.create_field("f")        # dense bernoulli rows (bitmap containers)
        ex.create_field("m")        # mixed array/bitmap/run containers
        ex.create_field("sp")       # sparse p=1e-4 (arrays / short runs)
        ex.create_field("z")        # zipf ranked field
        ex.create_field("exists")   # existence field for Not()
        ex.create_field("v", type="int", min=-(1 << 31), max=(1 << 31) - 1)
        ex.create_field("w", type="int", min=0, max=1000)
    for s in shards:
        frs = {
            "f": synth.bernoulli(s, 0, 6, 0.10),
            "m": synth.mixed(s, 0, 8),
            "sp": synth.bernoulli(s, 0, 4, 1e-4),
            "z": synth.zipf(s, n_rows=300, k=4),
            "exists": synth.bernoulli(s, 0, 1, 0.9),
            "v": synth.bsi(s, bit_depth=32, density=0.7),
            "w": synth.bsi(s, bit_depth=10, vmin=0, vmax=1000, density=0.01),
        }
        for name, fb in frs.items():
            for ex in (g, c):
                ex.load_fragment(name, s, fb)
    yield g, c, shards + [3]  # query shard 3 too: no fragment anywhere
    eng.close()


CALLS = [
    Intersect(Row("f", 1), Row("f", 2)),                                   # C1: bitmap x bitmap
    Intersect(Row("f", 1), Row("f", 2), Row("f", 3)),
    Union(Row("m", 0), Row("m", 1), Row("m", 2), Row("m", 3)),             # C2
    Difference(Row("m", 0), Row("m", 1), Row("m", 2), Row("m", 3)),
    Xor(Row("m", 0), Row("m", 1), Row("m", 2), Row("m", 3)),
    Intersect(Row("sp", 1), Row("sp", 2)),                                  # C5: sparse
    Intersect(Row("sp", 0), Row("m", 5)),
    Intersect(Row("m", 4), Row("f", 0)),
    Union(Row("sp", 0), Row("sp", 1), Row("sp", 2), Row("sp", 3)),
    Row("m", 6),
    Row("f", 99),                                                           # missing row
    Intersect(Row("f", 1), Row("f", 99)),
    Union(Row("f", 99), Row("m", 1)),
    Difference(Row("m", 1), Row("f", 99)),
    Intersect(Union(Row("m", 0), Row("f", 1)), Union(Row("m", 2), Row("sp", 3))),      # bushy: stack
    Difference(Row("f", 0), Union(Row("m", 1), Row("m", 2))),
    Xor(Intersect(Row("m", 0), Row("m", 1)), Difference(Row("m", 2), Row("m", 3)), Row("f", 5)),
    Union(Intersect(Row("m", 0), Row("m", 1)), Intersect(Row("m", 2), Row("m", 3)), Intersect(Row("m", 4), Row("m", 5))),
    Not(Row("m", 1)),
    Not(Intersect(Row("f", 1), Row("m", 2))),
    Union(),
    Xor(),
    Empty(),
    Intersect(Row("m", 7)),
]


@pytest.mark.parametrize("idx", range(len(CALLS)))
def test_count_matches_oracle(pair, idx):
    g, c, shards = pair
    call = CALLS[idx]
    assert g.count(call, shards) == c.count(call, shards), repr(call)


def test_count_per_shard_and_batch(pair):
    g, c, shards = pair
    for call in CALLS[:8]:
        assert g.be.count_shards(g.program(call), shards) == c.be.count_shards(c.program(call), shards)
    got = g.count_batch(CALLS, shards)
    exp = [c.count(call, shards) for call in CALLS]
    assert got == exp


def test_count_edge_shard_lists(pair):
    g, c, _ = pair
    call = Intersect(Row("f", 1), Row("f", 2))
    assert g.count(call, []) == 0
    assert g.count(call, [3]) == 0            # no fragment
    assert g.count(call, [10 ** 6]) == 0      # beyond the family table
    assert g.count(call, [2, 2]) == 2 * g.count(call, [2])  # duplicates are independent map jobs


def test_random_trees(pair):
    g, c, shards = pair
    rnd = random.Random(7)
    leaves = [("f", r) for r in range(6)] + [("m", r) for r in range(8)] + [("sp", r) for r in range(4)] + [("f", 77)]

    def tree(depth):
        if depth == 0 or rnd.random() < 0.3:
            f, r = rnd.choice(leaves)
            return Row(f, r)
        op = rnd.choice([Intersect, Union, Difference, Xor])
        return op(*[tree(depth - 1) for _ in range(rnd.randint(1, 3))])

    n_ok = 0
    for _ in range(60):
        t = tree(3)
        try:
            got = g.count(t, shards)
        except Exception as e:  # only the documented limits may be hit
            assert "too" in str(e), str(e)
            continue
        assert got == c.count(t, shards), repr(t)
        n_ok += 1
    assert n_ok > 40


def test_eval_columns_and_roaring(pair):
    g, c, shards = pair
    for call in [CALLS[0], CALLS[2], CALLS[3], CALLS[4], CALLS[5], CALLS[14], CALLS[18]]:
        for s in (0, 5, 3):
            gp, cp = g.program(call), c.program(call)
            assert g.be.eval_columns(gp, s) == c.be.eval_columns(cp, s), (repr(call), s)
            assert g.be.eval_roaring(gp, s) == c.be.eval_roaring(cp, s), (repr(call), s)


def test_row_segments_serialised_on_device(pair):
    """pb200_eval_shards (K11): every shard's result typed by Container.optimize and serialised by the
    device kernels must equal, byte for byte, the oracle's serialisation of the reference result."""
    g, c, shards = pair
    ex = Row("exists", 0)
    calls = [CALLS[i] for i in (0, 2, 3, 4, 5, 6, 8, 9, 10, 14, 16, 18, 20, 22)]
    calls += [Shift(ex, n=1), Shift(Union(Row("m", 1), ex), n=3), Xor(Shift(ex, n=2), Row("sp", 1)), Not(Row("exists", 0)),
              g.row_range("v", abi.BSI_GT, 0), g.row_range("w", abi.BSI_BETWEEN, 10, 500), Intersect(Row("sp", 0), g.row_range("v", abi.BSI_LT, -5))]
    kinds = set()
    for call in calls:
        got = g.row_segments(call, shards + [5, 0])   # duplicates: independent map jobs
        exp = c.row_segments(call, shards + [5, 0])
        assert got == exp, repr(call)
        for _, b in got:
            keyn = int.from_bytes(b[4:8], "little")
            kinds |= {int.from_bytes(b[8 + 12 * k + 8: 8 + 12 * k + 10], "little") for k in range(keyn)}
    assert kinds == {1, 2, 3}  # array, bitmap and run containers all came out of the device path
    assert g.row_segments(calls[0], []) == []


def test_shift_with_carry_out_of_the_shard(pair):
    """Shift (executor.go:1771, row.go:217, roaring.go:944): the bit leaving column 2^20-1 stays in the
    shard's segment under the next shard's container key and takes part in counts, set ops and the
    serialized row exactly as in the reference (several of these rows end in a set last column)."""
    g, c, shards = pair
    ex, m1, f1 = Row("exists", 0), Row("m", 1), Row("f", 1)
    dense = Union(*[Row("f", r) for r in range(6)], ex)
    calls = [Shift(ex, n=1), Shift(ex, n=2), Shift(dense, n=1), Shift(dense, n=5), Shift(Shift(dense, n=1), n=2), Shift(m1, n=1),
             Intersect(Shift(ex, n=1), Shift(dense, n=2)), Union(Shift(ex, n=1), f1), Intersect(Shift(ex, n=1), f1),
             Xor(Shift(ex, n=3), Shift(dense, n=1)), Difference(Shift(dense, n=2), Shift(ex, n=1)), Difference(f1, Shift(dense, n=1)),
             Shift(Intersect(Shift(ex, n=1), Shift(dense, n=1)), n=1), Not(Shift(ex, n=1)), Shift(Row("f", 99), n=1), Shift(Empty(), n=1),
             Union(Intersect(Shift(ex, n=1), Shift(dense, n=1)), Intersect(Shift(m1, n=2), Shift(dense, n=3)))]
    carried = 0
    for call in calls:
        gp, cp = g.program(call), c.program(call)
        assert g.count(call, shards) == c.count(call, shards), repr(call)
        assert g.be.count_shards(gp, shards) == c.be.count_shards(cp, shards), repr(call)
        for s in (0, 5, 3):
            cols = g.be.eval_columns(gp, s)
            assert cols == c.be.eval_columns(cp, s), (repr(call), s)
            assert g.be.eval_roaring(gp, s) == c.be.eval_roaring(cp, s), (repr(call), s)
            carried += sum(1 for x in cols if x >= (s + 1) << 20)
    assert carried > 10  # the case under test really occurs
    assert g.count_batch(calls + CALLS[:4], shards) == [c.count(call, shards) for call in calls + CALLS[:4]]
    # as a TopN source / BSI filter only the shard's own columns can match: same as the reference, where
    # the candidate rows and bit planes have no container under the next shard's key
    src = Shift(dense, n=1)
    assert g.topn("z", shards, n=10, src=src) == c.topn("z", shards, n=10, src=src)
    assert g.sum("v", shards, src) == c.sum("v", shards, src)
    assert g.min("v", shards, src) == c.min("v", shards, src)


def test_rows(pair):
    g, c, shards = pair
    col = (5 << 20) + 12345
    for fld in ("f", "m", "sp", "z", "exists"):
        for kw in (dict(), dict(limit=3), dict(previous=2), dict(previous=1, limit=2), dict(column=777), dict(column=col), dict(column=col, limit=5),
                   dict(previous=10 ** 9), dict(limit=0)):
            assert g.rows(fld, shards, **kw) == c.rows(fld, shards, **kw), (fld, kw)


def test_group_by(pair):
    """GroupBy (executor.go:1071-1275, :3060-3230): every combination counted on the device, iterator rules
    replayed on the host, against the oracle's literal restatement of the nested iterator."""
    from pilosa_b200.executor import RowsCall as R

    g, c, shards = pair
    flt = Intersect(Row("exists", 0), Union(Row("f", 1), Row("m", 2)))
    cases = [
        ([R("f")], {}), ([R("f")], dict(filter_call=flt)), ([R("m")], dict(filter_call=Row("sp", 1), limit=3)),
        ([R("f"), R("m")], {}), ([R("f"), R("m")], dict(filter_call=flt)), ([R("f"), R("m")], dict(limit=7)),
        ([R("m"), R("sp")], dict(filter_call=Row("f", 0))), ([R("sp"), R("sp")], {}),
        ([R("z", limit=12), R("f")], dict(limit=40)), ([R("f"), R("z", previous=100, limit=9)], {}),
        ([R("f", previous=2), R("m", previous=3)], dict(limit=10)), ([R("f", previous=5), R("m", previous=7)], {}),
        ([R("f", previous=77), R("m")], {}), ([R("m", column=(2 << 20) + 5), R("f")], {}),
        ([R("f"), R("m"), R("sp")], {}), ([R("f"), R("m"), R("sp")], dict(filter_call=flt, limit=25)),
        ([R("f", previous=1), R("m", previous=7), R("sp", previous=3)], dict(limit=6)),
        ([R("f", previous=1), R("m", previous=2), R("sp", previous=1)], dict(filter_call=Row("exists", 0))),
        ([R("sp"), R("f"), R("m"), R("sp")], dict(limit=50)),
        ([R("f"), R("m")], dict(filter_call=g.row_range("v", abi.BSI_GT, 0))),
        ([R("f"), R("m")], dict(filter_call=Row("f", 99))),      # empty filter
        ([R("f"), R("w")], {}),                                   # int field: its standard view has no fragments
    ]
    nonempty = 0
    for children, kw in cases:
        fams_ok = all(ch.field in g.fields for ch in children)
        if not fams_ok:
            continue
        got = g.group_by(children, shards, **kw)
        assert got == c.group_by(children, shards, **kw), (children, kw)
        nonempty += bool(got)
    assert nonempty >= 15


def test_topn(pair):
    g, c, shards = pair
    src = Intersect(Row("f", 1), Row("f", 2))
    for kw in (dict(n=10), dict(n=10, src=src), dict(n=50, src=Row("m", 1)), dict(n=0, src=src), dict(n=5, src=src, threshold=30),
               dict(n=3, src=Row("sp", 1)), dict(n=10, src=src, ids=[1, 5, 7, 250, 9999]), dict(n=10, ids=[3, 4, 299]),
               dict(n=5, src=src, tanimoto=1), dict(n=5, src=Row("z", 3), tanimoto=30)):
        assert g.topn("z", shards, **kw) == c.topn("z", shards, **kw), kw
    # single shard, exact fragment.top order (heap pop order)
    for s in (0, 5):
        gp, cp = g.program(src), c.program(src)
        assert g.be.topn_shard(g.family("z"), gp, 7, s) == c.be.topn_shard(c.family("z"), cp, 7, s)
        assert g.be.topn_shard(g.family("z"), None, 7, s) == c.be.topn_shard(c.family("z"), None, 7, s)


def test_bsi_sum_min_max(pair):
    g, c, shards = pair
    filt = Row("f", 1)
    filt2 = Intersect(Row("f", 1), Row("m", 2))
    for fld in ("v", "w"):
        for f in (None, filt, filt2, Row("f", 99)):
            assert g.sum(fld, shards, f) == c.sum(fld, shards, f), (fld, repr(f))
            assert g.min(fld, shards, f) == c.min(fld, shards, f), (fld, repr(f))
            assert g.max(fld, shards, f) == c.max(fld, shards, f), (fld, repr(f))
    # strict flag (opt-in deviation from the reference quirk) agrees with the oracle's strict mode
    fam, prog = g.family("v", "bsig_v"), g.program(filt)
    assert g.be.bsi_sum(fam, 32, prog, shards, flags=abi.SUM_STRICT) == c.be.bsi_sum(c.family("v", "bsig_v"), 32, c.program(filt), shards, flags=abi.SUM_STRICT)


def test_bsi_range(pair):
    g, c, shards = pair
    ops = [abi.BSI_EQ, abi.BSI_NEQ, abi.BSI_LT, abi.BSI_LTE, abi.BSI_GT, abi.BSI_GTE]
    for x in (0, 1, -1, 12345, -12345, 1 << 30, -(1 << 30), (1 << 31) - 1, -(1 << 31), 5 * 10 ** 9):
        for op in ops:
            assert g.count(g.row_range("v", op, x), shards) == c.count(c.row_range("v", op, x), shards), (op, x)
    for lo, hi in ((0, 100), (-100, 100), (-5000, -10), (-(1 << 31), (1 << 31) - 1), (10, 5), (1 << 29, 1 << 30)):
        assert g.count(g.row_range("v", abi.BSI_BETWEEN, lo, hi), shards) == c.count(c.row_range("v", abi.BSI_BETWEEN, lo, hi), shards)
    assert g.count(g.row_range("v", abi.BSI_NEQ, None), shards) == c.count(c.row_range("v", abi.BSI_NEQ, None), shards)
    # sparse small-depth field: array/run planes exercise the staged plane loads
    for x in (0, 1, 500, 999, 1000, 1001, -1):
        for op in ops:
            assert g.count(g.row_range("w", op, x), shards) == c.count(c.row_range("w", op, x), shards), ("w", op, x)
    # range leaf inside a larger tree, and materialised
    call = Intersect(Row("f", 1), g.row_range("v", abi.BSI_GT, 0))
    ccall = Intersect(Row("f", 1), c.row_range("v", abi.BSI_GT, 0))
    assert g.count(call, shards) == c.count(ccall, shards)
    assert g.be.eval_columns(g.program(call), 1) == c.be.eval_columns(c.program(ccall), 1)


def test_frag_apply_and_reload(pair):
    g, c, shards = pair
    fam_g, fam_c = g.family("sp"), c.family("sp")
    rnd = random.Random(3)
    sets = [r * (1 << 20) + rnd.randrange(1 << 20) for r in (0, 1, 1, 2, 9) for _ in range(40)]
    clears = sets[::5] + [3 * (1 << 20) + 17]
    g.be.frag_apply(fam_g, 1, sets, clears)
    for p in sets:
        c.be.set_bit(fam_c, 1, p >> 20, p & ((1 << 20) - 1) | (1 << 20))
    for p in clears:
        c.be.clear_bit(fam_c, 1, p >> 20, p & ((1 << 20) - 1) | (1 << 20))
    for call in (Row("sp", 1), Intersect(Row("sp", 1), Row("sp", 2)), Row("sp", 9), Union(Row("sp", 0), Row("sp", 9))):
        assert g.count(call, [1]) == c.count(call, [1])
        assert g.be.eval_columns(g.program(call), 1) == c.be.eval_columns(c.program(call), 1)


def test_import_roaring(pair):
    """pb200_frag_import_roaring = fragment.importRoaring (fragment.go:2243) -> ImportRoaringBits (roaring.go:1463):
    official-format and Pilosa-format payloads, set and clear, changed counts, new shard, then queries agree."""
    from test_abi_and_host import _official_roaring

    g, c, shards = pair
    rnd = random.Random(11)
    words = [rnd.getrandbits(64) & rnd.getrandbits(64) for _ in range(1024)]
    # fragment positions row * 2^20 + col: container key = row * 16 + (col >> 16); rows 2 and 40 of field "m"
    official = _official_roaring([(2 * 16 + 1, "a", sorted(rnd.sample(range(65536), 900))), (2 * 16 + 3, "b", words),
                                  (40 * 16 + 0, "a", [0, 1, 65535])])
    with_runs = _official_roaring([(2 * 16 + 1, "r", [(10, 5000), (60000, 65535)]), (5 * 16 + 15, "a", [3, 9])], with_runs=True)
    pilosa_fmt = synth.mixed(0, 2, 2).bytes()  # rows 2..3 in Pilosa format
    fm_g, fm_c = g.family("m"), c.family("m")
    for shard, blob, clear in ((0, official, False), (0, with_runs, False), (1, pilosa_fmt, False), (0, with_runs, True),
                               (2, official, True), (9, official, False), (0, official, False)):
        assert g.be.frag_import_roaring(fm_g, shard, blob, clear) == c.be.frag_import_roaring(fm_c, shard, blob, clear), (shard, clear)
    sh = shards + [9]
    for call in (Row("m", 2), Row("m", 40), Union(Row("m", 2), Row("m", 5)), Xor(Row("m", 2), Row("m", 3), Row("f", 1))):
        assert g.count(call, sh) == c.count(call, sh), repr(call)
        assert g.row_segments(call, sh) == c.row_segments(call, sh), repr(call)
    assert g.rows("m", sh) == c.rows("m", sh)
    from pilosa_b200.engine import EngineError
    with pytest.raises(EngineError, match="serialCookie"):
        g.be.frag_import_roaring(fm_g, 0, b"\x01\x02\x03\x04\x05\x06\x07\x08\x09")
    # undo, so that later tests see the original fixture
    for ex in (g, c):
        for s_ in (0, 1, 2):
            ex.load_fragment("m", s_, synth.mixed(s_, 0, 8))
        ex.be.frag_drop(ex.family("m"), 9)


def test_import_roaring_rank_recount():
    """The rank-cache update of fragment.importRoaring (fragment.go:2258-2276) on the engine's own rankCache state
    (cache.go:136-288): BulkAdd(row, Get(row) + changes) + Recalculate after every import, with a cache of 4 entries so
    that the admission threshold, the 1.1 x trim of the entries, a count that stays stale because 0 is below the
    threshold, and the uint64 wrap of a clear on an uncached row all happen. After every step the ranking
    (pb200_rank_get = rankCache.Top()) and TopN with and without a source row equal the oracle's."""
    from pilosa_b200.engine import Engine

    def frag(rows):  # {row: [cols]} -> Pilosa roaring bytes of shard 0
        return oracle.Bitmap(*sorted(r * (1 << 20) + c for r, cols in rows.items() for c in cols)).serialize()

    eng, orc = Engine(0), oracle.Oracle(workers=2)
    try:
        g, c = Executor(eng), Executor(orc)
        for ex in (g, c):
            ex.create_field("t")
            ex.create_field("s")
        fg, fc = g.family("t"), c.family("t")
        eng.set_option(Engine.OPT_RANK_CACHE_SIZE, 4)
        orc.cache_config(fc, 0, False, 4)
        rnd = random.Random(5)
        base = {r: sorted(rnd.sample(range(70000), 10 + 7 * r)) for r in range(8)}  # counts 10, 17, .. 59: rows 4..7 are cached
        src = frag({0: list(range(0, 70000, 3))})
        for ex in (g, c):
            ex.load_fragment("t", 0, frag(base))
            ex.load_fragment("s", 0, src)

        def check(step):
            assert eng.rank_get(fg, 0) == orc.rankings(fc, 0), step
            for n in (0, 2, 3):
                assert g.topn("t", [0], n=n) == c.topn("t", [0], n=n), (step, n)
                assert g.topn("t", [0], n=n, src=Row("s", 0)) == c.topn("t", [0], n=n, src=Row("s", 0)), (step, n)

        check("load")
        assert eng.rank_get(fg, 0)[0] == [7, 6, 5, 4]
        steps = [
            ("cached row grows", {5: list(range(100000, 100020))}, False),
            ("uncached row, below the threshold: dropped", {1: list(range(200000, 200005))}, False),
            ("uncached row, above the threshold: enters with the CHANGE as its count", {2: list(range(300000, 300060))}, False),
            ("new row", {30: list(range(0, 400, 2))}, False),
            ("top row cleared completely: 0 is below the threshold, the old count stays", {30: list(range(0, 400, 2))}, True),
            ("clear on an uncached row: Get() = 0 minus the change wraps", {0: base[0][:3]}, True),
            ("bits that are already set: nothing changes", {7: base[7][:5]}, False),
            ("partial clear of a cached row", {6: base[6][:20]}, True),
        ]
        for name, rows, clear in steps:
            blob = frag(rows)
            assert eng.frag_import_roaring(fg, 0, blob, clear) == orc.frag_import_roaring(fc, 0, blob, clear), name
            check(name)
        ids, counts = eng.rank_get(fg, 0)
        assert counts[0] == (1 << 64) - 3 and ids[0] == 0  # the wrapped count leads the ranking, as it would in the reference
        # pb200_frag_apply = importPositions (fragment.go:2051-2098): the rows the positions name are re-read (:2080-2093)
        SWd = 1 << 20
        pos_steps = [
            ("uncached row re-read: enters with its real count", [2 * SWd + 5000], []),
            ("row emptied: 0 is below the threshold, the cached count stays", [], [r_ * SWd + c_ for r_, c_ in ((0, c_) for c_ in base[0][3:])]),
            ("two rows named, one loses its wrapped count", [1 * SWd + 9000], [6 * SWd + base[6][20]]),
            ("a bit that is already set still names its row", [7 * SWd + base[7][0]], []),
            ("new row by positions", [77 * SWd + c_ for c_ in range(0, 900, 3)], []),
        ]
        for name, set_pos, clear_pos in pos_steps:
            eng.frag_apply(fg, 0, set_pos, clear_pos)
            orc.frag_import_positions(fc, 0, set_pos, clear_pos)
            check(name)
            assert g.count(Union(*[Row("t", r_) for r_ in range(8)]), [0]) == c.count(Union(*[Row("t", r_) for r_ in range(8)]), [0])
        # a list handed over by the host replaces the engine's state
        eng.rank_set(fg, 0, [3, 4], [9, 8])
        assert eng.rank_get(fg, 0) == ([3, 4], [9, 8])
    finally:
        eng.close()


def _row_message_class():
    """internal.Row { repeated uint64 Columns = 1; } (internal/public.proto) built with google.protobuf: an independent
    protobuf implementation to check the bytes against."""
    from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

    fdp = descriptor_pb2.FileDescriptorProto(name="public_min.proto", package="internal", syntax="proto3")
    m = fdp.message_type.add(name="Row")
    m.field.add(name="Columns", number=1, type=descriptor_pb2.FieldDescriptorProto.TYPE_UINT64,
                label=descriptor_pb2.FieldDescriptorProto.LABEL_REPEATED)
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fdp)
    return message_factory.GetMessageClass(pool.FindMessageTypeByName("internal.Row"))


def test_proto_row_columns():
    """pb200_eval_proto_columns: the Columns field of internal.Row sized and written on the device (K12) equals the
    oracle's restatement of the generated marshaller (public.pb.go:745-766) and google.protobuf's encoding of the same
    column list: shard 0 (1-, 2- and 3-byte varints inside one container), shards on both sides of 2^21 and 2^28, a
    dense and a sparse row, set operations, a BSI range, Shift with its carry into the next shard, an empty row."""
    from pilosa_b200.engine import Engine

    RowMsg = _row_message_class()
    eng, orc = Engine(0), oracle.Oracle(workers=2)
    try:
        g, c = Executor(eng), Executor(orc)
        shards = [0, 1, 2, 3, 255, 256, 70000]
        for ex in (g, c):
            ex.create_field("m")
            ex.create_field("sp")
            ex.create_field("v", type="int", min=-1000, max=1000)
            for s in shards:
                ex.load_fragment("m", s, synth.mixed(s, 0, 6))
                ex.load_fragment("sp", s, synth.bernoulli(s, 0, 2, 2e-4))
                ex.load_fragment("v", s, synth.bsi(s, 10, -1000, 1000, 0.02))
        low = oracle.Bitmap(*([0, 1, 5, 127, 128, 129, 300, 16383, 16384, 16385, 65535, 65536, (1 << 20) - 1] + [(1 << 20) + x for x in (0, 127, 128)])).serialize()
        for ex in (g, c):  # row 0 / row 1 of field "lo" in shard 0: the varint-length steps inside the first container
            ex.create_field("lo")
            ex.load_fragment("lo", 0, low)
        calls = [Row("lo", 0), Union(Row("lo", 0), Row("lo", 1)), Row("sp", 0), Row("m", 1), Row("m", 3), Intersect(Row("m", 0), Row("m", 2)),
                 Xor(Row("m", 1), Row("sp", 1)), Difference(Row("m", 4), Row("m", 5)), Row("sp", 1),
                 g.row_range("v", abi.BSI_GT, 900), g.row_range("v", abi.BSI_BETWEEN, -5, 5),
                 Shift(Row("m", 2), n=1), Shift(Row("sp", 0), n=3), Row("m", 99), Intersect(Row("sp", 0), Row("sp", 1))]
        for call in calls:
            for sh in (shards, [0], [255, 256], [3, 1], [70000]):
                got = g.row_proto_columns(call, sh)
                assert got == c.row_proto_columns(call, sh), (repr(call), sh)
                cols = [x for s in sorted(sh) for x in g.be.eval_columns(g.program(call), s)]
                assert got == RowMsg(Columns=cols).SerializeToString(), (repr(call), sh)
        assert g.row_proto_columns(Row("m", 99), shards) == b""  # no columns: the field is omitted
        assert g.row_proto_columns(Row("lo", 0), [0])[:2] == bytes([0x0A, len(g.row_proto_columns(Row("lo", 0), [0])) - 2])
        import ctypes

        p = g.program(Row("m", 1))
        n = ctypes.c_size_t()
        buf = ctypes.create_string_buffer(4)
        rc = eng.lib.pb200_eval_proto_columns(eng._h, p, len(p), abi.u64_array([0]), 1, buf, 4, ctypes.byref(n))
        assert rc == -6 and n.value == len(g.row_proto_columns(Row("m", 1), [0]))  # PB200_ERR_RANGE carries the needed size
    finally:
        eng.close()


def test_densify_option_changes_layout_not_results():
    """PB200_OPT_DENSIFY_MIN_BYTES: large array / run containers kept as bitmaps in HBM; every result and the
    algorithmic byte count stay what they are without it."""
    from pilosa_b200.engine import Engine

    shards = [0, 1, 2]
    calls = [Union(*[Row("m", r) for r in range(4)]), Difference(*[Row("m", r) for r in range(4)]), Xor(Row("m", 0), Row("m", 5), Row("sp", 1)),
             Intersect(Row("m", 1), Row("sp", 0)), Not(Row("m", 2)), Shift(Row("m", 3), n=1)]
    res = []
    for thr in (0, 512):
        eng = Engine(0)
        if thr:
            eng.set_option(Engine.OPT_DENSIFY_MIN_BYTES, thr)
        ex = Executor(eng)
        for name in ("m", "sp", "z", "exists"):
            ex.create_field(name)
        for s in shards:
            ex.load_fragment("m", s, synth.mixed(s, 0, 8))
            ex.load_fragment("sp", s, synth.bernoulli(s, 0, 4, 1e-4))
            ex.load_fragment("z", s, synth.zipf(s, n_rows=200, k=4))
            ex.load_fragment("exists", s, synth.bernoulli(s, 0, 1, 0.9))
        out = [ex.count(c_, shards) for c_ in calls]
        algo = eng.last_timing().algo_bytes
        out.append([ex.row_segments(c_, shards) for c_ in calls[:3]])
        out.append(ex.topn("z", shards, n=10, src=Row("m", 1)))
        out.append(ex.group_by([__import__("pilosa_b200.executor", fromlist=["RowsCall"]).RowsCall("m")], shards, filter_call=Row("z", 3)))
        eng.frag_apply(ex.family("m"), 0, set_pos=[5, (1 << 20) * 3 + 77], clear_pos=[(1 << 20) + 9])  # rewrite goes through the option too
        out.append(ex.count(calls[0], shards))
        res.append((out, algo, eng.frag_stat(ex.family("m"), 0), eng.frag_stat(ex.family("sp"), 0)))
        eng.close()
    (plain, algo0, m0, sp0), (dense, algo1, m1, sp1) = res
    assert plain == dense
    assert algo0 == algo1                      # byte accounting follows the reference's encoding
    assert m1[2] > m0[2] and sp1[2] == sp0[2]  # payload bytes: mixed field grew, the tiny arrays stayed


def test_time_range_rows():
    """Row(f=r, from=, to=) on a time field (executor.go:1501-1531): the union over the quantum views, engine vs oracle."""
    from datetime import datetime

    from pilosa_b200.engine import Engine

    eng = Engine(0)
    g, c = Executor(eng), Executor(oracle.Oracle(workers=4))
    shards = [0, 1, 3]
    views = ["standard_2000112822", "standard_2000112823", "standard_20001129", "standard_200012", "standard_2001", "standard_200201",
             "standard_2002030100", "standard_1999", "standard"]
    for ex in (g, c):
        ex.create_field("t", type="time", time_quantum="YMDH")
        ex.create_field("f")
    for s in shards:
        for i, v in enumerate(views):
            if s == 3 and i % 3 == 0:
                continue  # some views have no fragment in some shards (:1519-1522)
            fb = synth.mixed(s, 0, 3, 100 + i) if i % 2 else synth.bernoulli(s, 0, 3, 0.02, 100 + i)
            for ex in (g, c):
                ex.load_fragment("t", s, fb, view=v)
        for ex in (g, c):
            ex.load_fragment("f", s, synth.bernoulli(s, 0, 2, 0.1))
    t = lambda s_: datetime.strptime(s_, "%Y-%m-%d %H:%M")
    ranges = [("2000-11-28 22:00", "2002-03-01 03:00"), ("2000-11-28 23:00", "2000-11-30 00:00"), ("2001-01-01 00:00", "2002-01-01 00:00"),
              ("1990-01-01 00:00", "1991-01-01 00:00"), ("2000-12-01 00:00", "2002-02-01 00:00")]
    seen = set()
    for a, b in ranges:
        for row in (0, 2):
            gc, cc = g.row_time("t", row, t(a), t(b)), c.row_time("t", row, t(a), t(b))
            assert repr(gc) == repr(cc)
            seen.add(gc.name)
            assert g.count(gc, shards) == c.count(cc, shards), (a, b, row)
            assert g.row_segments(gc, shards) == c.row_segments(cc, shards), (a, b, row)
            both = Intersect(gc, Row("f", 1))
            assert g.count(both, shards) == c.count(Intersect(cc, Row("f", 1)), shards)
    assert seen == {"Union", "Row", "Empty"}
    eng.close()


def test_wide_nary_operations():
    """The reference puts no limit on the operands of Union / Intersect / Xor / Difference (executor.go:1667-1735) and a
    time-range Row is a Union over as many views as the range has (:1501-1531). A call whose leaves do not fit one
    device program is compiled with its wide operations evaluated chunk-wise into dense temp rows; every result must
    equal the oracle's, on the count, batch, per-shard, serialised-row and protobuf paths."""
    from datetime import datetime

    from pilosa_b200.engine import Engine

    eng = Engine(0)
    try:
        g, c = Executor(eng), Executor(oracle.Oracle(workers=4))
        shards = [0, 1, 2]
        for ex in (g, c):
            ex.create_field("w")   # 100 sparse rows
            ex.create_field("d")   # 48 dense rows
            ex.create_field("m")   # mixed containers
            ex.create_field("t", type="time", time_quantum="H")
            for s in shards:
                ex.load_fragment("w", s, synth.bernoulli(s, 0, 100, 0.004))
                ex.load_fragment("d", s, synth.bernoulli(s, 0, 48, 0.9))
                ex.load_fragment("m", s, synth.mixed(s, 0, 8))
        hours = [f"standard_200011{d:02d}{h:02d}" for d in (27, 28) for h in range(24)]  # 48 hourly views
        for i, v in enumerate(hours):
            for s in shards:
                if (i + s) % 7 == 0:
                    continue
                fb = synth.bernoulli(s, 0, 2, 0.003, 500 + i)
                for ex in (g, c):
                    ex.load_fragment("t", s, fb, view=v)
        W = lambda a, b: [Row("w", r) for r in range(a, b)]  # noqa: E731
        D = lambda a, b: [Row("d", r) for r in range(a, b)]  # noqa: E731
        calls = [
            Union(*W(0, 100)), Xor(*W(0, 70)), Intersect(*D(0, 40)), Intersect(*D(0, 33)), Union(*W(0, 33)),
            Difference(Row("m", 1), *W(0, 50)), Difference(Row("d", 0), *W(0, 100)),
            Intersect(Row("m", 2), Union(*W(0, 50)), Xor(*W(50, 90))),
            Union(Shift(Row("m", 3), n=1), *W(0, 60)),                      # a Shift operand compiled before the wide run
            Union(Intersect(Row("m", 0), Row("m", 1)), *W(10, 75)),         # leading non-leaf operand stays an operand
            Xor(Union(*W(0, 40)), Union(*W(30, 80)), Row("m", 4)),
            Intersect(Union(*W(0, 100)), Row("d", 3)),
            Difference(Union(*W(0, 64)), Union(*W(32, 96))),
        ]
        for call in calls:
            assert g.count(call, shards) == c.count(call, shards), repr(call)[:80]
            assert g.be.count_shards(g.program(call), shards) == c.be.count_shards(c.program(call), shards), repr(call)[:80]
        assert g.count_batch(calls, shards) == c.count_batch(calls, shards)
        assert g.count_batch(calls[:3] + [Row("m", 1)] + calls[7:9], shards) == c.count_batch(calls[:3] + [Row("m", 1)] + calls[7:9], shards)
        for call in (calls[0], calls[5], calls[7], calls[8]):
            assert g.row_segments(call, shards) == c.row_segments(call, shards), repr(call)[:80]
            assert g.row_proto_columns(call, shards) == c.row_proto_columns(call, shards), repr(call)[:80]
            assert g.row_columns(call, [1]) == c.row_columns(call, [1])
        # a time range over 2 days of hourly views: 48 Row leaves in one Union
        t0, t1 = datetime(2000, 11, 27, 0), datetime(2000, 11, 29, 0)
        tr = g.row_time("t", 1, t0, t1)
        assert repr(tr) == repr(c.row_time("t", 1, t0, t1)) and len(tr.children) > 32
        assert g.count(tr, shards) == c.count(tr, shards) > 0
        both = Intersect(tr, Row("d", 2))
        assert g.count(both, shards) == c.count(both, shards)
        assert g.row_segments(both, shards) == c.row_segments(both, shards)
    finally:
        eng.close()


def test_ops_log_replay_on_load(pair):
    g, c, _ = pair
    base = synth.bernoulli(7, 0, 2, 0.001).bytes()
    ops = oracle.encode_op(0, [5]) + oracle.encode_op(2, [1 << 20, (1 << 20) + 70000, 900000]) + oracle.encode_op(1, [5]) + \
        oracle.encode_op(3, [900000, 12])
    for ex in (g, c):
        ex.create_field("ol")
        ex.load_fragment("ol", 7, base + ops)
    for call in (Row("ol", 0), Row("ol", 1), Intersect(Row("ol", 0), Row("ol", 1))):
        assert g.be.eval_columns(g.program(call), 7) == c.be.eval_columns(c.program(call), 7)


def test_error_behaviour(pair):
    from pilosa_b200.engine import EngineError

    g, _, shards = pair
    with pytest.raises(EngineError, match="empty Intersect"):
        g.count(Intersect(), shards)
    with pytest.raises(EngineError, match="empty Difference"):
        g.count(Difference(), shards)
    with pytest.raises(EngineError, match="too many"):  # every step adds a temp-row leaf: the 32-leaf limit comes first
        g.count(Shift(Row("f", 1), n=65), shards)
    assert g.count(Shift(Row("f", 1), n=31), shards) > 0
    with pytest.raises(EngineError):
        g.be.frag_load(g.family("f"), 0, b"\x00\x01\x02")
    with pytest.raises(EngineError, match="magic"):
        g.be.frag_load(g.family("f"), 0, b"\x00" * 64)
    # a failed load leaves the resident fragment intact
    assert g.count(Row("f", 1), [0]) > 0


def test_plain_c_host_runs():
    """The C example (no Python, no torch in the process) drives load / count / eval_shards / TopN through the C ABI."""
    import os
    import subprocess

    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "examples", "count_c")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "C host ok" in r.stdout, r.stdout + r.stderr
