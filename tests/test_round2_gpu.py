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


def _cont(kind, vals):
    """(type, n, payload) of a container holding `vals` (sorted) in the given encoding."""
    vals = sorted(vals)
    if kind == "a":
        return 1, len(vals), struct.pack("<%dH" % len(vals), *vals)
    if kind == "b":
        words = [0] * 1024
        for v in vals:
            words[v >> 6] |= 1 << (v & 63)
        return 2, len(vals), struct.pack("<1024Q", *words)
    runs, st, prev = [], None, None
    for v in vals:
        if st is None:
            st = prev = v
        elif v == prev + 1:
            prev = v
        else:
            runs.append((st, prev))
            st = prev = v
    runs.append((st, prev))
    return 3, len(vals), struct.pack("<H", len(runs)) + b"".join(struct.pack("<HH", a, b) for a, b in runs)


def test_fold_kernel_edge_patterns(eng_orc):
    """The encoding-native fold evaluator (eval_fold_kernel) on hand-made containers: run lists with the maximum of 2048
    runs, runs touching 0 and 65535, values packed into one word or spread one per word, arrays of 1 / 4095 / 4096
    values — every ordered pair under all four operations, and left folds of three and four operands, against the oracle."""
    import itertools

    eng, orc = eng_orc
    g, c = Executor(eng), Executor(orc)
    for ex in (g, c):
        ex.create_field("e")
    rnd = __import__("random").Random(5)
    pats = [
        ("a", [7]),                                               # one value
        ("a", list(range(64, 96))),                               # one word full of values, array-encoded
        ("a", list(range(0, 65536, 16))[:4095]),                  # 4095 values, one or two per word
        ("a", sorted(rnd.sample(range(65536), 4096))),            # ArrayMaxSize values, unoptimized
        ("a", [0, 31, 32, 63, 65504, 65535]),                     # word and container boundaries
        ("r", [v for k in range(2048) for v in (32 * k, 32 * k + 1)]),              # 2048 runs: the run maximum
        ("r", [v for k in range(2048) for v in range(32 * k + 30, 32 * k + 32)]),   # 2048 runs ending on word boundaries, last at 65535
        ("r", list(range(0, 40000))),                             # one long run from 0
        ("r", list(range(100, 65536))),                           # one long run to 65535
        ("r", list(range(5, 9)) + list(range(40, 2000)) + list(range(2001, 2003)) + list(range(60000, 65000))),
        ("r", [v for k in range(300) for v in range(200 * k, 200 * k + 150)]),       # 300 medium runs
        ("b", sorted(rnd.sample(range(65536), 30000))),
        ("b", sorted(rnd.sample(range(65536), 5000))),
        # slices of 8 values / 4 runs that share words with their neighbours in every possible way
        ("a", list(range(0, 4096))),                               # 32 values per word: four slices per word
        ("a", list(range(1000, 1000 + 37)) + list(range(3000, 3000 + 91)) + [65535]),
        ("a", [v for k in range(400) for v in range(160 * k + 27, 160 * k + 37)]),   # clusters of 10 straddling word borders
        ("r", list(range(0, 4096, 2))),                            # 2048 runs of one value: 16 starts and 16 ends per word
        ("r", [v for k in range(1500) for v in range(40 * k + 29, 40 * k + 35)]),    # 1500 runs, every one across a word border
        ("r", [v for k in range(1023) for v in range(64 * k + 3, 64 * k + 64)] + list(range(65480, 65536))),  # 1024 runs, each ends on a word border
        ("R", [(10 * k, 10 * k + 9) for k in range(1000)] + [(20000, 20000), (20001, 65535)]),  # adjacent runs, not merged (a reader accepts them)
    ]
    cs = []
    for row, (kind, vals) in enumerate(pats):
        if kind == "R":   # a raw run list, written as given
            n = sum(b - a + 1 for a, b in vals)
            t, pl = 3, struct.pack("<H", len(vals)) + b"".join(struct.pack("<HH", a, b) for a, b in vals)
            pats[row] = ("r", [v for a, b in vals for v in range(a, b + 1)])
        else:
            t, n, pl = _cont(kind, vals)
        cs.append((row * 16 + 5, t, n, pl))
        if row % 3 == 0:   # a second column so that items differ inside one query
            t2, n2, pl2 = _cont("b", sorted(rnd.sample(range(65536), 20000)))
            cs.append((row * 16 + 9, t2, n2, pl2))
    cs.sort()
    data = pilosa_file(cs)
    for ex in (g, c):
        ex.load_fragment("e", 0, data)
    sets = [set(v) for _, v in pats]
    R = lambda i: Row("e", i)  # noqa: E731
    calls, want = [], []
    ops = [(Intersect, lambda a, b: a & b), (abi.Union, lambda a, b: a | b), (abi.Xor, lambda a, b: a ^ b), (abi.Difference, lambda a, b: a - b)]
    for i, j in itertools.permutations(range(len(pats)), 2):
        for op, _ in ops:
            calls.append(op(R(i), R(j)))
    for trip in ((0, 5, 11), (5, 2, 8), (8, 6, 3), (11, 5, 6), (3, 9, 10), (6, 7, 12), (9, 3, 5), (2, 10, 7), (12, 6, 9), (7, 8, 5),
                 (13, 16, 19), (16, 14, 17), (19, 15, 18), (17, 18, 13), (15, 19, 16), (14, 13, 12)):
        for op, _ in ops:
            calls.append(op(*[R(i) for i in trip]))
            calls.append(op(R(trip[0]), R(trip[1]), R(trip[2]), R(trip[1]), R(4)))
    got = []
    for k in range(0, len(calls), 256):
        got += g.count_batch(calls[k:k + 256], [0])
    exp = [c.count(x, [0]) for x in calls]
    bad = [(repr(calls[k]), got[k], exp[k]) for k in range(len(calls)) if got[k] != exp[k]]
    assert not bad, bad[:5]
    # column 5 alone, from the sets (pins the oracle on these patterns as well)
    for (i, j), (op, f) in itertools.product(((2, 5), (5, 6), (3, 9), (7, 8), (13, 16), (15, 17), (19, 14), (18, 16)), ops):
        cols = [v for v in g.row_columns(op(R(i), R(j)), [0]) if (v >> 16) == 5]
        assert [v & 0xffff for v in cols] == sorted(f(sets[i], sets[j])), (i, j, op)


def test_writes_on_the_resident_copy_interleaved_with_queries(eng_orc):
    """pb200_frag_apply / pb200_frag_import_roaring rewrite only the touched containers of the resident fragment (a kernel
    per write, nothing of the fragment crosses PCIe): single bits, batches, new rows inside and far outside a dense row
    table (the table is rebuilt on the device, dense -> sparse), rows emptied again, imports that add and subtract, all
    interleaved with queries and compared with the oracle after every step."""
    import random
    import time

    eng, orc = eng_orc
    g, c = Executor(eng), Executor(orc)
    for ex in (g, c):
        ex.create_field("f")
        ex.create_field("m")
    shards = [0, 1]
    for s in shards:
        fb, mb = synth.bernoulli(s, 0, 64, 0.10), synth.mixed(s, 0, 8)
        for ex in (g, c):
            ex.load_fragment("f", s, fb)
            ex.load_fragment("m", s, mb)
    ff, fm = g.family("f"), g.family("m")
    cf, cm = c.family("f"), c.family("m")
    rnd = random.Random(17)

    def check(rows_f=(0, 1, 2, 63, 64, 70, 5000), rows_m=(0, 1, 2, 3, 9)):
        calls = [Row("f", r) for r in rows_f] + [Row("m", r) for r in rows_m]
        calls += [Intersect(Row("f", 1), Row("f", 70)), abi.Union(Row("m", 0), Row("m", 9), Row("f", 5000)), abi.Xor(Row("m", 1), Row("m", 2), Row("f", 2)),
                  abi.Difference(Row("f", 0), Row("m", 3), Row("f", 64))]
        assert g.count_batch(calls, shards) == [c.count(x, shards) for x in calls]
        for fld in ("f", "m"):
            assert g.rows(fld, shards) == c.rows(fld, shards), fld
        for call in (Row("f", 70), Row("m", 9), abi.Union(Row("m", 0), Row("m", 1))):
            assert g.row_segments(call, shards) == c.row_segments(call, shards), repr(call)
        assert g.topn("m", shards, n=5) == c.topn("m", shards, n=5)
        assert g.topn("f", shards, n=4, src=Row("m", 1)) == c.topn("f", shards, n=4, src=Row("m", 1))

    def both_apply(fg, fc, shard, sets, clears):
        eng.frag_apply(fg, shard, sets, clears)
        orc.frag_import_positions(fc, shard, sets, clears)

    check()
    # one bit, timed and metered: a write must move bytes, not the fragment (64 rows x 128 KiB here)
    pos = 1 * SW + 12345
    both_apply(ff, cf, 0, [pos], [])
    t0 = time.perf_counter()
    both_apply(ff, cf, 0, [2 * SW + 777], [])
    dt = time.perf_counter() - t0
    tm = eng.last_timing()
    assert tm.h2d_bytes + tm.d2h_bytes < 65536, (tm.h2d_bytes, tm.d2h_bytes)
    assert dt < 0.05
    check()
    # new rows: inside the dense table's slack, then far outside (the table turns sparse), then in between
    both_apply(ff, cf, 0, [64 * SW + 5, 64 * SW + 6, 70 * SW + 99999], [])
    check()
    both_apply(ff, cf, 1, [5000 * SW + 1, 5000 * SW + 70000, 70 * SW + 3], [1 * SW + 3])
    check()
    both_apply(ff, cf, 1, [300 * SW + k * 17 for k in range(5000)], [])     # an array container grows into a bitmap / runs
    both_apply(ff, cf, 1, [], [300 * SW + k * 17 for k in range(0, 5000, 2)])
    check(rows_f=(0, 1, 70, 300, 5000))
    # a row emptied again, and clears on rows / containers that do not exist
    both_apply(ff, cf, 0, [], [64 * SW + 5, 64 * SW + 6, 9999 * SW + 1, 3 * SW + 5])
    check()
    assert 64 not in g.rows("f", [0])
    # mixed containers: random batches of sets and clears over existing containers of every type
    for step in range(4):
        sets = [rnd.randrange(10) * SW + rnd.randrange(SW) for _ in range(400)]
        clears = [rnd.randrange(10) * SW + rnd.randrange(SW) for _ in range(400)]
        both_apply(fm, cm, step % 2, sets, clears)
        check()
    # imports: union and subtraction of serialized bitmaps, Pilosa format, into existing and new rows
    imp = synth.mixed(0, 7, 4).bytes()     # rows 7..10 of the generator
    for clear in (False, True, False):
        assert eng.frag_import_roaring(fm, 0, imp, clear) == orc.frag_import_roaring(cm, 0, imp, clear)
        check()
    assert eng.rank_get(fm, 0) == orc.rankings(cm, 0)
    # a reload replaces the written copy wholesale
    fb = synth.bernoulli(0, 0, 64, 0.10)
    for ex in (g, c):
        ex.load_fragment("f", 0, fb)
    check()


def test_sixteen_host_threads_on_one_context():
    """examples/concurrent_c: 16 OS threads x 1000 single pb200_count calls on ONE context; every answer is checked
    against the single-threaded one inside the program (which exits non-zero on any difference), and the combining
    queue must lift the aggregate rate well above the one-call-at-a-time rate."""
    import json
    import os
    import subprocess

    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "examples", "concurrent_c")
    out = subprocess.run([exe, "16", "1000"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    rec = json.loads(out.stdout.strip().splitlines()[-1])
    assert rec["speedup"] >= 3.0, rec


def test_concurrent_counts_match_oracle(eng_orc):
    """The same from Python threads with DIFFERENT programs and shard lists in flight together (ctypes drops the GIL
    inside the call): requests with equal shard lists are combined into one batch, the others run in their own rounds,
    a malformed program fails alone — every result equals the oracle's."""
    import threading

    eng, orc = eng_orc
    g, c = Executor(eng), Executor(orc)
    for ex in (g, c):
        ex.create_field("f")
        ex.create_field("m")
    shards = [0, 1, 2]
    for s in shards:
        fb, mb = synth.bernoulli(s, 0, 8, 0.10), synth.mixed(s, 0, 6)
        for ex in (g, c):
            ex.load_fragment("f", s, fb)
            ex.load_fragment("m", s, mb)
    calls = [Intersect(Row("f", a), Row("f", b)) for a in range(8) for b in range(a)] + [abi.Union(Row("m", a), Row("m", (a + 1) % 6), Row("f", a)) for a in range(6)]
    lists = [shards, [0], [2, 1], shards]
    want = {(i, j): c.count(call, lists[j]) for i, call in enumerate(calls) for j in range(len(lists))}
    progs = [g.program(x) for x in calls]
    errors, bad_seen = [], []

    def worker(tid):
        for k in range(150):
            i, j = (tid * 31 + k * 7) % len(calls), (tid + k) % len(lists)
            try:
                got = eng.count(progs[i], lists[j])
                if got != want[(i, j)]:
                    errors.append((tid, i, j, got, want[(i, j)]))
            except Exception as e:  # noqa: BLE001
                errors.append((tid, i, j, repr(e)))
            if k % 50 == 7:
                try:
                    eng.count((abi.Node * 1)(abi.Node(op=abi.OP_INTERSECT, nargs=0)), lists[j])
                    errors.append((tid, "malformed program accepted"))
                except Exception:  # noqa: BLE001
                    bad_seen.append(tid)

    th = [threading.Thread(target=worker, args=(t,)) for t in range(12)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errors, errors[:5]
    assert len(bad_seen) == 12 * 3


def test_topn_pruned_like_fragment_top(eng_orc):
    """TopN with a source counts only the ranked rows fragment.top's walk visits (fragment.go:1673-1688): same pairs as
    the oracle and as the engine with pruning switched off, for several n / thresholds, with fewer bytes read."""
    eng, orc = eng_orc
    g, c = Executor(eng), Executor(orc)
    for ex in (g, c):
        ex.create_field("z")
        ex.create_field("a")
    shards = [0, 1, 2, 3]
    for s in shards:
        zb, ab = synth.zipf(s, n_rows=1500, k=6), synth.bernoulli(s, 1, 2, 0.10)
        for ex in (g, c):
            ex.load_fragment("z", s, zb)
            ex.load_fragment("a", s, ab)
    src = Intersect(Row("a", 1), Row("a", 2))
    cases = [dict(n=5, src=src), dict(n=50, src=src), dict(n=1, src=Row("a", 1)), dict(n=20, src=src, threshold=40), dict(n=400, src=src),
             dict(n=2000, src=Row("a", 2)), dict(n=10, src=Row("z", 3)), dict(n=7, src=abi.Union(Row("z", 0), Row("a", 1)))]
    OPT_TOPN_PRUNE = 3
    for kw in cases:
        want = c.topn("z", shards, **kw)
        eng.set_option(OPT_TOPN_PRUNE, 1)
        got = g.topn("z", shards, **kw)
        pruned_bytes = eng.last_timing().algo_bytes
        eng.set_option(OPT_TOPN_PRUNE, 0)
        full = g.topn("z", shards, **kw)
        full_bytes = eng.last_timing().algo_bytes
        eng.set_option(OPT_TOPN_PRUNE, 1)
        assert got == want, kw
        assert full == want, kw
        if kw["n"] <= 5:  # (a pruned pass stages the source twice: on this small index it only pays when the walk stops early)
            assert pruned_bytes < full_bytes, (kw, pruned_bytes, full_bytes)
    # per shard, exactly fragment.top's list and order
    for s in shards:
        assert g.be.topn_shard(g.family("z"), g.program(src), 9, s) == c.be.topn_shard(c.family("z"), c.program(src), 9, s)
    # a host-supplied ranking that is NOT in descending order of the cached counts: the walk's rule, not sortedness, decides
    ids, counts = orc.rankings(c.family("z"), 0)
    ids2, counts2 = ids[:40][::-1] + ids[40:], counts[:40][::-1] + counts[40:]
    eng.rank_set(g.family("z"), 0, ids2, counts2)
    orc_pairs = None
    try:
        orc.rank_set(c.family("z"), 0, ids2, counts2)
        orc_pairs = c.be.topn_shard(c.family("z"), c.program(src), 6, 0)
    except AttributeError:
        pass
    got = g.be.topn_shard(g.family("z"), g.program(src), 6, 0)
    eng.set_option(OPT_TOPN_PRUNE, 0)
    assert got == g.be.topn_shard(g.family("z"), g.program(src), 6, 0)
    eng.set_option(OPT_TOPN_PRUNE, 1)
    if orc_pairs is not None:
        assert got == orc_pairs


def test_reserve_then_counts_do_not_allocate(eng_orc):
    """pb200_reserve sizes the scratch once; the Count calls within its bounds then leave the device's free memory
    untouched (no cudaMalloc / cudaFree on the query path) and still agree with the oracle."""
    import torch

    from pilosa_b200.abi import Difference, Union, Xor

    eng, orc = eng_orc
    g, c = Executor(eng), Executor(orc)
    shards = list(range(24))
    for ex in (g, c):
        ex.create_field("m")
        for s in shards:
            ex.load_fragment("m", s, synth.mixed(s, 0, 4, 7))
    rows = [Row("m", r) for r in range(4)]
    calls = [Union(*rows), Difference(*rows), Xor(*rows), Intersect(*rows), Intersect(Union(rows[0], rows[1]), Xor(rows[2], rows[3]))]
    eng.reserve(64, len(shards), 7)
    g.count(rows[0], shards)                       # first call: lazily created streams / events, if any are left
    torch.cuda.synchronize(0)
    free0 = torch.cuda.mem_get_info(0)[0]
    for call in calls:
        assert g.count(call, shards) == c.count(call, shards), repr(call)
    batch = [calls[i % len(calls)] for i in range(64)]
    assert g.count_batch(batch, shards) == c.count_batch(batch, shards)
    torch.cuda.synchronize(0)
    assert torch.cuda.mem_get_info(0)[0] == free0
    with pytest.raises(Exception):
        eng.reserve(0, 4, 3)


def test_lane_per_item_pass_on_very_sparse_rows(eng_orc):
    """eval_tiny_kernel: rows of a few values per container (p = 1e-4 .. 3e-4) — pairs of tiny arrays are counted one lane
    per item, longer arrays / runs / three-operand programs are handed on to the other evaluators; totals, per-shard
    counts and batches against the oracle, with shards missing from one field."""
    from pilosa_b200.abi import Difference, Union, Xor

    eng, orc = eng_orc
    g, c = Executor(eng), Executor(orc)
    shards = list(range(700))   # x 16 containers: more items than the device holds warps, or the pass is not used for a single call
    for ex in (g, c):
        ex.create_field("sp")
        ex.create_field("md")
        for s in shards:
            ex.load_fragment("sp", s, synth.bernoulli(s, 1, 4, 1e-4, 11))
            if s % 7 != 3:
                ex.load_fragment("md", s, synth.bernoulli(s, 1, 3, 3e-4, 12))     # ~20 values: many containers beyond 16
    A, B, C_, D = Row("sp", 1), Row("sp", 2), Row("sp", 3), Row("sp", 4)
    M, N = Row("md", 1), Row("md", 2)
    calls = [A, M, Row("sp", 99)]
    for op in (Intersect, Union, Xor, Difference):
        calls += [op(A, B), op(B, A), op(A, A), op(A, M), op(M, N), op(N, A), op(A, Row("sp", 99)), op(Row("md", 77), B), op(A, B, C_), op(M, A, D, N)]
    for call in calls:
        assert g.count(call, shards) == c.count(call, shards), repr(call)
        assert g.be.count_shards(g.program(call), shards) == c.be.count_shards(c.program(call), shards), repr(call)
    assert g.count_batch(calls * 3, shards[:60]) == c.count_batch(calls * 3, shards[:60])
    assert g.count_batch(calls, shards[5:6]) == c.count_batch(calls, shards[5:6])
    # self-consistency on every shard: |A| + |B| = |A or B| + |A and B|
    n = {k: g.count(x, shards) for k, x in dict(a=A, b=B, u=Union(A, B), i=Intersect(A, B), x=Xor(A, B), d=Difference(A, B)).items()}
    assert n["a"] + n["b"] == n["u"] + n["i"] and n["x"] == n["u"] - n["i"] and n["d"] == n["a"] - n["i"]


def test_proto_columns_of_short_containers_as_value_lists(eng_orc):
    """K12 over results whose containers hold 0, 1, 31, 32 (kept as 64-byte value lists) and 33, 40, 5000 values (kept
    as words), side by side in one shard and across shards, for plain rows and for folds: the Columns field against the
    oracle's, byte for byte."""
    from pilosa_b200.abi import Difference, Union, Xor

    eng, orc = eng_orc
    g, c = Executor(eng), Executor(orc)
    rnd = __import__("random").Random(9)
    sizes = [0, 1, 31, 32, 33, 40, 5000, 2, 32, 33, 7, 0, 31, 64, 3, 65536]
    for ex in (g, c):
        ex.create_field("l")
    for shard in (0, 3, 4, 9):
        cs = []
        for row in range(3):
            for col, n in enumerate(sizes):
                n = (n + row * 5) % 65537 if n not in (0, 65536) else n
                if n == 0:
                    continue
                vals = list(range(65536)) if n == 65536 else sorted(rnd.sample(range(65536), n))
                t, nn, pl = _cont("a" if n <= 4096 else "b", vals)
                cs.append((row * 16 + col, t, nn, pl))
        data = pilosa_file(sorted(cs))
        for ex in (g, c):
            ex.load_fragment("l", shard, data)
    shards = [0, 3, 4, 5, 9]
    R = lambda i: Row("l", i)  # noqa: E731
    for call in (R(0), R(1), R(2), R(7), Intersect(R(0), R(1)), Union(R(0), R(1), R(2)), Xor(R(0), R(2)), Difference(R(1), R(0))):
        assert g.row_proto_columns(call, shards) == c.be.eval_proto_columns(c.program(call), shards), repr(call)
        assert g.row_proto_columns(call, shards[1:2]) == c.be.eval_proto_columns(c.program(call), shards[1:2]), repr(call)
        assert g.row_columns(call, shards) == c.row_columns(call, shards), repr(call)


def test_fold_kernel_random_clustered_containers(eng_orc):
    """Randomised stress of the fold evaluator's slice ownership and run fill: containers whose values cluster inside
    single words and straddle slice borders (8 values / 4 runs per thread), in all three encodings, folded with all
    four operations over 2..5 operands; every count against Python sets (and the oracle on a sample)."""
    import random

    from pilosa_b200.abi import Difference, Union, Xor

    eng, orc = eng_orc
    g, c = Executor(eng), Executor(orc)
    rnd = random.Random(20261107)

    def clustered(n_clusters, max_len, step_choices):
        vals, pos = set(), 0
        for _ in range(n_clusters):
            pos += rnd.choice(step_choices)
            ln = rnd.randint(1, max_len)
            for k in range(ln):
                if pos + k < 65536 and rnd.random() < 0.8:
                    vals.add(pos + k)
            pos += ln
            if pos >= 65536:
                break
        return sorted(vals)

    n_rows, cols = 48, (0, 7, 15)
    sets, cs = {}, []
    for row in range(n_rows):
        for col in cols:
            kind = rnd.choice("aarrb")
            if kind == "a":
                vals = clustered(rnd.randint(1, 500), rnd.choice((3, 9, 40)), (1, 2, 5, 31, 33, 64, 300))[:4096]
            elif kind == "r":
                vals = clustered(rnd.randint(1, 1500), rnd.choice((2, 6, 70, 400)), (2, 3, 7, 32, 65, 500))
            else:
                vals = sorted(rnd.sample(range(65536), rnd.randint(4097, 40000)))
            if not vals:
                vals = [rnd.randrange(65536)]
            if kind == "r":   # keep the run encoding legal: at most 2048 runs
                runs = 1 + sum(1 for i in range(1, len(vals)) if vals[i] != vals[i - 1] + 1)
                if runs > 2048:
                    kind = "a" if len(vals) <= 4096 else "b"
            if kind == "a" and len(vals) > 4096:
                kind = "b"
            t, n, pl = _cont(kind, vals)
            cs.append((row * 16 + col, t, n, pl))
            sets[(row, col)] = set(vals)
    data = pilosa_file(sorted(cs))
    for ex in (g, c):
        ex.create_field("k")
        ex.load_fragment("k", 0, data)
    ops = [(Intersect, lambda a, b: a & b), (Union, lambda a, b: a | b), (Xor, lambda a, b: a ^ b), (Difference, lambda a, b: a - b)]
    calls, want = [], []
    for _ in range(600):
        op, f = rnd.choice(ops)
        rows = [rnd.randrange(n_rows) for _ in range(rnd.randint(2, 5))]
        calls.append(op(*[Row("k", r) for r in rows]))
        tot = 0
        for col in cols:
            acc = set(sets[(rows[0], col)])
            for r in rows[1:]:
                acc = f(acc, sets[(r, col)])
            tot += len(acc)
        want.append(tot)
    got = []
    for k in range(0, len(calls), 200):
        got += g.count_batch(calls[k:k + 200], [0])
    bad = [(repr(calls[i]), got[i], want[i]) for i in range(len(calls)) if got[i] != want[i]]
    assert not bad, bad[:3]
    for i in range(0, len(calls), 37):
        assert c.count(calls[i], [0]) == want[i]
        assert g.row_columns(calls[i], [0]) == c.row_columns(calls[i], [0])
