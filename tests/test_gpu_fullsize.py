"""Parity at BASELINE sizes (one test per config family), in the driver's GPU run: the engine holds the config's full
shard count; the oracle holds a sample of the shards (it would need minutes for all of them) and must agree, shard
by shard, with the engine's per-shard answers there; over ALL shards the size-independent identities of the domain
are checked (inclusion-exclusion, partitions of a range, per-shard sums = totals)."""
import os
from concurrent.futures import ThreadPoolExecutor

import pytest

import oracle
from pilosa_b200 import abi, synth
from pilosa_b200.abi import Difference, Intersect, Row, Union, Xor
from pilosa_b200.executor import Executor

pytestmark = pytest.mark.gpu
SEED = 20190416


def _load(g, c, field, gen, shards, sample):
    with ThreadPoolExecutor(max_workers=min(16, os.cpu_count() or 8)) as tp:
        for s, fb in zip(shards, tp.map(gen, shards)):
            g.load_fragment(field, s, fb)
            if s in sample:
                c.load_fragment(field, s, fb)


@pytest.fixture()
def pair():
    from pilosa_b200.engine import Engine

    eng = Engine(0)
    yield Executor(eng), Executor(oracle.Oracle(workers=min(16, os.cpu_count() or 8)))
    eng.close()


def test_c2_mixed_containers_1024_shards(pair):
    """BASELINE configs[1]: Count(Union / Difference / Xor) over 4 rows of mixed array / bitmap / run containers, 1024 shards."""
    g, c = pair
    shards = list(range(1024))
    sample = set(range(0, 1024, 64)) | {1023}
    for ex in (g, c):
        ex.create_field("m")
    _load(g, c, "m", lambda s: synth.mixed(s, 0, 4, SEED), shards, sample)
    rows = [Row("m", r) for r in range(4)]
    calls = [Union(*rows), Difference(*rows), Xor(*rows), Intersect(rows[0], rows[1]), Intersect(*rows)]
    ss = sorted(sample)
    for call in calls:
        per = g.be.count_shards(g.program(call), shards)
        assert [per[s] for s in ss] == c.be.count_shards(c.program(call), ss), repr(call)
        assert sum(per) == g.count(call, shards)
    a, b = rows[0], rows[1]
    n = {k: g.count(x, shards) for k, x in dict(a=a, b=b, u=Union(a, b), i=Intersect(a, b), x=Xor(a, b), d=Difference(a, b)).items()}
    assert n["a"] + n["b"] == n["u"] + n["i"] and n["x"] == n["u"] - n["i"] and n["d"] == n["a"] - n["i"]


def test_c4_bsi_256_shards(pair):
    """BASELINE configs[3] at 256 of its 2048 shards (1 GiB of planes): Sum / Min / Max / Range, 32-bit values."""
    g, c = pair
    shards = list(range(256))
    sample = set(range(0, 256, 32)) | {255}
    for ex in (g, c):
        ex.create_field("v", type="int", min=-(1 << 31), max=(1 << 31) - 1)
        ex.create_field("fl")
    _load(g, c, "v", lambda s: synth.bsi(s, 32, -(1 << 31), (1 << 31) - 1, 1.0, SEED), shards, sample)
    _load(g, c, "fl", lambda s: synth.bernoulli(s, 1, 1, 0.10, SEED), shards, sample)
    ss = sorted(sample)
    for filt in (None, Row("fl", 1)):
        assert g.sum("v", ss, filt) == c.sum("v", ss, filt)
        assert g.min("v", ss, filt) == c.min("v", ss, filt)
        assert g.max("v", ss, filt) == c.max("v", ss, filt)
    for x in (0, 1 << 30, -(1 << 30), 12345):
        call_g, call_c = g.row_range("v", abi.BSI_GT, x), c.row_range("v", abi.BSI_GT, x)
        per = g.be.count_shards(g.program(call_g), shards)
        assert [per[s] for s in ss] == c.be.count_shards(c.program(call_c), ss), x
        le = g.count(g.row_range("v", abi.BSI_LTE, x), shards)
        assert sum(per) + le == 256 << 20  # every column holds a value
    (mn, _), (mx, _) = g.min("v", shards), g.max("v", shards)
    assert g.count(g.row_range("v", abi.BSI_LT, mn), shards) == 0 and g.count(g.row_range("v", abi.BSI_GT, mx), shards) == 0
    fam, bd = g.family("v", "bsig_v"), 32
    s_all, c_all = g.sum("v", shards)
    s_f, c_f = g.be.bsi_sum(fam, bd, g.program(Row("fl", 1)), shards, flags=abi.SUM_STRICT)
    s_n, c_n = g.be.bsi_sum(fam, bd, g.program(Difference(g.row_range("v", abi.BSI_NEQ, None), Row("fl", 1))), shards, flags=abi.SUM_STRICT)
    assert s_f + s_n == s_all and c_f + c_n == c_all == 256 << 20


def test_c5_sparse_8192_shards(pair):
    """BASELINE configs[4] at its largest size: p = 1e-4 rows (arrays / short runs), 8192 shards."""
    g, c = pair
    shards = list(range(8192))
    sample = set(range(0, 8192, 512)) | {8191}
    for ex in (g, c):
        ex.create_field("sp")
    _load(g, c, "sp", lambda s: synth.bernoulli(s, 1, 2, 1e-4, SEED), shards, sample)
    ss = sorted(sample)
    for call in (Intersect(Row("sp", 1), Row("sp", 2)), Union(Row("sp", 1), Row("sp", 2)), Row("sp", 1)):
        per = g.be.count_shards(g.program(call), shards)
        assert [per[s] for s in ss] == c.be.count_shards(c.program(call), ss), repr(call)
        assert sum(per) == g.count(call, shards)
    assert g.row_proto_columns(Row("sp", 2), ss) == c.be.eval_proto_columns(c.program(Row("sp", 2)), ss)


def test_c3_topn_256_shards(pair):
    """BASELINE configs[2] at 256 of its 1024 shards, 10 000 rows: TopN(n=50) with an Intersect source; the counts of
    the returned ids are re-derived with explicit Count(Intersect(src, Row)) calls over all shards."""
    g, c = pair
    shards = list(range(256))
    sample = {0, 100, 255}
    for ex in (g, c):
        ex.create_field("z")
        ex.create_field("a")
    _load(g, c, "z", lambda s: synth.zipf(s, 10000, 8, 1.6, 50.0, SEED), shards, sample)
    _load(g, c, "a", lambda s: synth.bernoulli(s, 1, 2, 0.10, SEED), shards, sample)
    src = Intersect(Row("a", 1), Row("a", 2))
    ss = sorted(sample)
    assert g.topn("z", ss, n=50, src=src) == c.topn("z", ss, n=50, src=src)
    for s in ss:
        assert g.be.topn_shard(g.family("z"), g.program(src), 50, s) == c.be.topn_shard(c.family("z"), c.program(src), 50, s)
    top = g.topn("z", shards, n=50, src=src)
    assert len(top) == 50 and all(top[i][1] >= top[i + 1][1] for i in range(49))
    assert [n for _, n in top] == g.count_batch([Intersect(src, Row("z", i)) for i, _ in top], shards)
