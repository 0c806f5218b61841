"""GPU parity tests added in round 2: prepared batches against writes that happen between prepare and launch,
containers beyond the descriptor's 13-bit length field, timing bookkeeping after a failed call. All through the
C ABI, compared with the CPU oracle (bit-exact)."""
import ctypes as C
import struct

import pytest

import oracle
from pilosa_b200 import abi, synth
from pilosa_b200.abi import Intersect, Row
from pilosa_b200.executor import Executor

pytestmark = pytest.mark.gpu

SW = 1 << 20


def pilosa_file(containers):
    """Pilosa roaring format (roaring.go:1044-1122): containers = [(key, type, n, payload bytes)]."""
    out = struct.pack("<II", 12348, len(containers))
    for key, typ, n, _ in containers:
        out += struct.pack("<QHH", key, typ, n - 1)
    off = 8 + 16 * len(containers)
    for _, _, _, pl in containers:
        out += struct.pack("<I", off)
        off += len(pl)
    for _, _, _, pl in containers:
        out += pl
    return out


def oversize_fragments():
    """Row 0: an ARRAY container of 10 000 values (> ArrayMaxSize); row 1: a RUN container of 9 000 runs (> runMaxSize);
    row 2: an ordinary array overlapping both. Un-optimized on purpose: the reference's reader accepts them."""
    arr = list(range(0, 40000, 4))                       # 10 000 values
    runs = [(7 * i, 7 * i + 2) for i in range(9000)]     # 9 000 runs of 3
    small = list(range(0, 60000, 12))
    cs = [
        (0 * 16 + 3, 1, len(arr), struct.pack("<%dH" % len(arr), *arr)),
        (1 * 16 + 3, 3, 3 * len(runs), struct.pack("<H", len(runs)) + b"".join(struct.pack("<HH", a, b) for a, b in runs)),
        (2 * 16 + 3, 1, len(small), struct.pack("<%dH" % len(small), *small)),
    ]
    return pilosa_file(cs), arr, runs, small


def official_runs_file(runs):
    """Official roaring format with the is-run bitmap (cookie 12347, roaring.go:5081-5245), one container, key 0."""
    n = sum(b - a + 1 for a, b in runs)
    out = struct.pack("<I", 12347 | (0 << 16))     # one container
    out += b"\x01"                                  # is-run bitmap
    out += struct.pack("<HH", 0, n - 1)
    out += struct.pack("<H", len(runs))
    out += b"".join(struct.pack("<HH", a, b - a) for a, b in runs)
    return out


@pytest.fixture()
def eng_orc():
    from pilosa_b200.engine import Engine

    eng = Engine(0)
    orc = oracle.Oracle(workers=2)
    yield eng, orc
    eng.close()


def test_oversize_containers_match_oracle(eng_orc):
    eng, orc = eng_orc
    g, c = Executor(eng), Executor(orc)
    for ex in (g, c):
        ex.create_field("o")
    data, arr, runs, small = oversize_fragments()
    for ex in (g, c):
        ex.load_fragment("o", 0, data)
    sa, sr, ss = set(arr), {v for a, b in runs for v in range(a, b + 1)}, set(small)
    calls = {
        "a": (Row("o", 0), len(sa)), "r": (Row("o", 1), len(sr)),
        "a&r": (Intersect(Row("o", 0), Row("o", 1)), len(sa & sr)),
        "a&s": (Intersect(Row("o", 0), Row("o", 2)), len(sa & ss)),
        "r&s": (Intersect(Row("o", 1), Row("o", 2)), len(sr & ss)),
        "a|r|s": (abi.Union(Row("o", 0), Row("o", 1), Row("o", 2)), len(sa | sr | ss)),
        "a^r^s": (abi.Xor(Row("o", 0), Row("o", 1), Row("o", 2)), len(sa ^ sr ^ ss)),
        "s-a-r": (abi.Difference(Row("o", 2), Row("o", 0), Row("o", 1)), len(ss - sa - sr)),
    }
    for name, (call, exp) in calls.items():
        assert c.count(call, [0]) == exp, name
        assert g.count(call, [0]) == exp, name
    # the materialised row too (serialised on the device, optimize rule applied)
    call = Intersect(Row("o", 0), Row("o", 1))
    assert g.row_columns(call, [0]) == sorted(3 * 65536 + v for v in sa & sr)


def test_oversize_official_import(eng_orc):
    eng, orc = eng_orc
    g, c = Executor(eng), Executor(orc)
    for ex in (g, c):
        ex.create_field("o")
    runs = [(6 * i, 6 * i + 1) for i in range(9000)]
    data = official_runs_file(runs)
    base = synth.bernoulli(0, 0, 2, 0.05)
    for ex in (g, c):
        ex.load_fragment("o", 0, base)
    n_c = orc.frag_import_roaring(c.family("o"), 0, data, False)
    n_g = eng.frag_import_roaring(g.family("o"), 0, data, False)
    assert n_g == n_c
    for call in (Row("o", 0), Row("o", 1), Intersect(Row("o", 0), Row("o", 1))):
        assert g.count(call, [0]) == c.count(call, [0])


def test_prepared_batch_survives_writes(eng_orc):
    """pb200_batch_prepare, then writes that put array containers into an all-bitmap family, grow the family's shard
    table and add a family, then pb200_batch_launch: the counts are those of the CURRENT fragments."""
    eng, orc = eng_orc
    g, c = Executor(eng), Executor(orc)
    for ex in (g, c):
        ex.create_field("f")
    shards = [0, 1]
    for s in shards:
        fb = synth.bernoulli(s, 0, 4, 0.10)
        for ex in (g, c):
            ex.load_fragment("f", s, fb)
    calls = [Intersect(Row("f", 0), Row("f", 1), Row("f", 9)), Intersect(Row("f", 0), Row("f", 1), Row("f", 2)), Intersect(Row("f", 9), Row("f", 2))]
    progs = [g.program(x) for x in calls]
    flat, off = eng.pack_programs(progs)
    sh = abi.u64_array(shards)
    b = eng.batch_prepare(flat, off, len(progs), sh, len(shards))
    out = (C.c_uint64 * len(progs))()
    eng.batch_launch(b)
    eng.batch_fetch(b, out)
    assert list(out) == [c.count(x, shards) for x in calls]
    # row 9 appears: a sparse array container per shard (the family stops being all-bitmap)
    pos = [9 * SW + 64 * k for k in range(0, 3000, 3)] + [9 * SW + 70000 + k for k in range(50)]
    for s in shards:
        eng.frag_apply(g.family("f"), s, pos, [])
        orc.frag_import_positions(c.family("f"), s, pos, [])
    # the shard table grows (shard 200 > initial capacity of 64) and a new family shifts the family table
    fb = synth.bernoulli(200, 0, 2, 0.10)
    for ex in (g, c):
        ex.load_fragment("f", 200, fb)
        ex.create_field("later")
    eng.batch_launch(b)
    eng.batch_fetch(b, out)
    exp = [c.count(x, shards) for x in calls]
    assert exp[0] > 0 and exp[2] > 0
    assert list(out) == exp
    assert list(out) == g.count_batch(calls, shards)
    eng.batch_free(b)


def test_timing_after_failed_count(eng_orc):
    """An invalid program must not leave the algorithmic-byte counter of the next call pointing at a batch slot."""
    eng, orc = eng_orc
    g = Executor(eng)
    g.create_field("v", type="int", min=0, max=1000)
    g.load_fragment("v", 0, synth.bsi(0, bit_depth=10, vmin=0, vmax=1000, density=0.5))
    fam = g.family("v", "bsig_v")
    eng.bsi_sum(fam, 10, None, [0])
    want = eng.last_timing().algo_bytes
    assert want > 0
    bad = (abi.Node * 1)(abi.Node(op=abi.OP_INTERSECT, nargs=0))
    with pytest.raises(Exception):
        eng.count(bad, [0])
    eng.bsi_sum(fam, 10, None, [0])
    assert eng.last_timing().algo_bytes == want
