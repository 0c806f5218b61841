"""Size-independent identities of the domain, checked on the CUDA engine alone (no oracle): the same checks
bench_configs.py runs at BASELINE.json's full sizes (1024 / 2048 shards), here on a small seeded index."""
import pytest

from pilosa_b200 import abi, synth
from pilosa_b200.abi import Difference, Intersect, Not, Row, Union, Xor
from pilosa_b200.executor import Executor

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def g():
    from pilosa_b200.engine import Engine

    eng = Engine(0)
    ex = Executor(eng)
    ex.create_field("m")
    ex.create_field("f")
    ex.create_field("z")
    ex.create_field("exists")
    ex.create_field("v", type="int", min=-(1 << 31), max=(1 << 31) - 1)
    for s in range(24):
        ex.load_fragment("m", s, synth.mixed(s, 0, 4))
        ex.load_fragment("f", s, synth.bernoulli(s, 0, 3, 0.10))
        ex.load_fragment("z", s, synth.zipf(s, n_rows=500, k=4))
        ex.load_fragment("exists", s, synth.bernoulli(s, 0, 1, 1.0))
        ex.load_fragment("v", s, synth.bsi(s, bit_depth=32, density=0.6))
    yield ex
    eng.close()


SH = list(range(24))


def test_inclusion_exclusion(g):
    rows = [Row("m", r) for r in range(4)] + [Row("f", 1)]
    for a in rows:
        for b in rows:
            ca, cb = g.count(a, SH), g.count(b, SH)
            ci, cu, cx, cd = (g.count(op(a, b), SH) for op in (Intersect, Union, Xor, Difference))
            assert ca + cb == cu + ci and cx == cu - ci and cd == ca - ci
    u = Union(*rows[:4])
    assert sum(g.be.count_shards(g.program(u), SH)) == g.count(u, SH)
    assert g.count(Not(u), SH) == g.count(Row("exists", 0), SH) - g.count(u, SH)  # every column exists
    assert g.count(Not(Not(u)), SH) == g.count(u, SH)


def test_bsi_partitions_and_extremes(g):
    rr = lambda op, *a: g.row_range("v", op, *a)
    nn = g.count(rr(abi.BSI_NEQ, None), SH)
    for x in (0, 1, -1, 12345, -(1 << 30), (1 << 31) - 1, -(1 << 31)):
        assert g.count(rr(abi.BSI_GT, x), SH) + g.count(rr(abi.BSI_LTE, x), SH) == nn
        assert g.count(rr(abi.BSI_LT, x), SH) + g.count(rr(abi.BSI_GTE, x), SH) == nn
        assert g.count(rr(abi.BSI_EQ, x), SH) + g.count(rr(abi.BSI_NEQ, x), SH) == nn
    (mn, cmn), (mx, cmx) = g.min("v", SH), g.max("v", SH)
    assert g.count(rr(abi.BSI_LT, mn), SH) == 0 and g.count(rr(abi.BSI_GT, mx), SH) == 0
    assert g.count(rr(abi.BSI_EQ, mn), SH) >= cmn >= 1 and g.count(rr(abi.BSI_EQ, mx), SH) >= cmx >= 1  # counts: first shard attaining it
    assert g.count(rr(abi.BSI_BETWEEN, -5000, 5000), SH) == g.count(rr(abi.BSI_LTE, 5000), SH) - g.count(rr(abi.BSI_LT, -5000), SH)
    fam = g.family("v", "bsig_v")
    s_all, c_all = g.be.bsi_sum(fam, 32, None, SH, flags=1)
    flt = Row("f", 1)
    s_f, c_f = g.be.bsi_sum(fam, 32, g.program(flt), SH, flags=1)
    s_n, c_n = g.be.bsi_sum(fam, 32, g.program(Difference(rr(abi.BSI_NEQ, None), flt)), SH, flags=1)
    assert (s_f + s_n, c_f + c_n) == (s_all, c_all) and c_all == nn
    pos, neg = g.be.bsi_sum(fam, 32, g.program(rr(abi.BSI_GT, 0)), SH, flags=1), g.be.bsi_sum(fam, 32, g.program(rr(abi.BSI_LT, 0)), SH, flags=1)
    assert pos[0] + neg[0] == s_all and pos[0] > 0 > neg[0]


def test_topn_counts_are_exact_and_stable(g):
    src = Intersect(Row("f", 1), Row("f", 2))
    top = g.topn("z", SH, n=25, src=src)
    assert [c for _, c in top] == g.count_batch([Intersect(src, Row("z", i)) for i, _ in top], SH)
    assert all(top[i][1] >= top[i + 1][1] for i in range(len(top) - 1))
    assert sorted(g.topn("z", SH, n=25, src=src, ids=[i for i, _ in top])) == sorted(top)
    from pilosa_b200.executor import RowsCall as R
    gb = dict(g.group_by([R("f"), R("m")], SH))
    for (a, b), c in gb.items():
        assert c == g.count(Intersect(Row("f", a), Row("m", b)), SH)
    assert sum(c for (a, _), c in g.group_by([R("f"), R("exists")], SH) if a == 1) == g.count(Row("f", 1), SH)


def test_row_segments_round_trip(g):
    """Serialised segments loaded back as a fragment give the same counts (encode -> decode round trip)."""
    call = Xor(Row("m", 0), Row("m", 1), Row("f", 2))
    segs = g.row_segments(call, SH)
    g.create_field("tmp")
    total = 0
    for s, data in segs:
        # a segment holds keys shard*16 + c: as a fragment of shard s that is row number `s` (fragment.go:629)
        g.load_fragment("tmp", s, data)
        total += g.count(Row("tmp", s), [s])
    assert total == g.count(call, SH)
