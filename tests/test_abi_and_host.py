"""CPU-side checks: the C-ABI library loads and exports every symbol the header declares,
the product refuses to run without a CUDA device (no fallback), and the host logic
(program flattening, BSI predicate re-basing, synthetic generator, file format)."""
import ctypes as C
import os
import subprocess
import sys

import pytest

import oracle
from pilosa_b200 import abi, synth
from pilosa_b200.abi import Difference, Intersect, Not, Row, Union, Xor
from pilosa_b200.executor import BsiGroup, Executor, bit_depth, bit_depth_int64

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from pilosa_b200 import engine

    lib = engine.load_library()
    syms = engine.declared_symbols()
    assert len(syms) >= 24
    for s in syms:
        assert hasattr(lib, s), s
    assert lib.pb200_abi_version() == 1
    assert lib.pb200_strerror(0) == b"ok"
    # the shared object itself (nm): every declared symbol is an exported text symbol
    out = subprocess.check_output(["nm", "-D", "--defined-only", engine.LIB_PATH]).decode()
    exported = {line.split()[-1] for line in out.splitlines() if " T " in line}
    assert set(syms) <= exported


def test_no_cpu_fallback():
    """Without a CUDA device pb200_init must fail loudly; with one it must succeed."""
    import torch

    from pilosa_b200.engine import Engine, EngineError

    if torch.cuda.is_available():
        Engine(0).close()
    else:
        with pytest.raises(EngineError, match="no CPU fallback"):
            Engine(0)


def test_product_never_imports_oracle():
    bad = []
    for dirpath, _, files in os.walk(os.path.join(ROOT, "pilosa_b200")):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".inl")):
                src = open(os.path.join(dirpath, fn)).read()
                for line in src.splitlines():
                    s = line.strip()
                    if fn == "selftest.py":
                        continue  # smoke(): the oracle is the checker there, by contract
                    if s.startswith(("import oracle", "from oracle")) or "oracle/" in s and "#include" in s:
                        bad.append((fn, s))
    assert not bad, bad


def test_flatten_post_order():
    fams = {}
    prog = abi.flatten(Difference(Row("a", 1), Union(Row("b", 2), Row("c", 3))), lambda f, v: fams.setdefault((f, v), len(fams)))
    ops = [(n.op, n.family, n.row, n.nargs) for n in prog]
    assert ops == [(abi.OP_ROW, 0, 1, 0), (abi.OP_ROW, 1, 2, 0), (abi.OP_ROW, 2, 3, 0), (abi.OP_UNION, 0, 0, 2), (abi.OP_DIFFERENCE, 0, 0, 2)]
    prog = abi.flatten(Not(Row("a", 1)), lambda f, v: 7)
    assert [(n.op, n.family) for n in prog] == [(abi.OP_ROW, 7), (abi.OP_NOT, 7)]


# field_internal_test.go:31-157 — baseValue / baseValueBetween tables, restated for Base -100/0/100
def test_bsi_group_base_value_tables():
    # the three groups and both tables are the literal rows of field_internal_test.go:31-157
    b0 = BsiGroup(min=-1000, max=1000, base=-100, bit_depth=10)
    b1 = BsiGroup(min=-255, max=255, base=0, bit_depth=8)
    b2 = BsiGroup(min=-(1 << 63), max=(1 << 63) - 1, base=100, bit_depth=11)
    LT, GT, EQ = abi.BSI_LT, abi.BSI_GT, abi.BSI_EQ
    cases = [
        (b0, LT, 5, 105, False), (b0, LT, -8, 92, False), (b0, LT, -108, -8, False), (b0, LT, 1005, 1023, False), (b0, LT, 0, 100, False),
        (b1, LT, 5, 5, False), (b1, LT, -8, -8, False), (b1, LT, 1005, 255, False), (b1, LT, 0, 0, False),
        (b2, LT, 5, -95, False), (b2, LT, -8, -108, False), (b2, LT, 105, 5, False), (b2, LT, 1105, 1005, False),
        (b0, GT, -5, 95, False), (b0, GT, 5, 105, False), (b0, GT, 905, 1005, False), (b0, GT, 0, 100, False),
        (b1, GT, 5, 5, False), (b1, GT, -8, -8, False), (b1, GT, 1005, 0, True), (b1, GT, 0, 0, False),
        (b2, GT, 5, -95, False), (b2, GT, -8, -108, False), (b2, GT, 105, 5, False), (b2, GT, 1105, 1005, False),
        (b0, EQ, -105, -5, False), (b0, EQ, 5, 105, False), (b0, EQ, 905, 1005, False), (b0, EQ, 0, 100, False),
        (b1, EQ, 5, 5, False), (b1, EQ, -8, -8, False), (b1, EQ, 1005, 0, True), (b1, EQ, 0, 0, False),
        (b2, EQ, 5, -95, False), (b2, EQ, -8, -108, False), (b2, EQ, 105, 5, False), (b2, EQ, 1105, 1005, False),
    ]
    for g, op, val, exp, oor in cases:
        assert g.base_value(op, val) == (exp, oor), (g.base, op, val)
        assert oracle.base_value(g.base, g.bit_depth, op, val) == (exp, oor)
    between = [
        (b0, -205, -105, -105, -5, False), (b0, -105, 80, -5, 180, False), (b0, 5, 20, 105, 120, False),
        (b0, 20, 1005, 120, 1023, False), (b0, 1005, 2000, 0, 0, True),
        (b1, -105, -5, -105, -5, False), (b1, -5, 20, -5, 20, False), (b1, 5, 20, 5, 20, False),
        (b1, 20, 1005, 20, 255, False), (b1, 1005, 2000, 0, 0, True),
        (b2, 5, 95, -95, -5, False), (b2, 95, 120, -5, 20, False), (b2, 105, 120, 5, 20, False),
        (b2, 120, 1105, 20, 1005, False), (b2, 1105, 2000, 1005, 1900, False),
    ]
    for g, lo, hi, elo, ehi, oor in between:
        assert g.base_value_between(lo, hi) == (elo, ehi, oor), (g.base, lo, hi)
        assert oracle.base_value_between(g.base, g.bit_depth, lo, hi) == (elo, ehi, oor)
    # field.go:1606-1621
    assert [bit_depth(v) for v in (0, 1, 2, 3, 4, 255, 256, (1 << 31) - 1, 1 << 31)] == [0, 1, 2, 2, 3, 8, 9, 31, 32]
    assert bit_depth_int64(-(1 << 31)) == 32 and oracle.lib().orc_bit_depth_i64(-(1 << 31)) == 32


def test_synth_is_seeded_and_canonical():
    a, b = synth.mixed(3, 0, 2).bytes(), synth.mixed(3, 0, 2).bytes()
    assert a == b
    assert synth.mixed(4, 0, 2).bytes() != a
    bm = oracle.Bitmap.deserialize(a)
    types = {t for _, t, _ in bm.containers()}
    assert types == {oracle.ARRAY, oracle.BITMAP, oracle.RUN}
    # canonical: re-serialising through the oracle's optimize() is the identity
    assert bm.serialize(optimize=True) == a
    dense = oracle.Bitmap.deserialize(synth.bernoulli(0, 1, 1, 0.10).bytes())
    assert all(t == oracle.BITMAP for _, t, _ in dense.containers()) and len(dense.containers()) == 16
    assert abs(dense.count() - 0.1 * (1 << 20)) < 3000
    sparse = oracle.Bitmap.deserialize(synth.bernoulli(0, 1, 1, 1e-4).bytes())
    assert all(t in (oracle.ARRAY, oracle.RUN) for _, t, _ in sparse.containers()) and 40 < sparse.count() < 200
    pos = [5, 70000, (3 << 20) + 9, (3 << 20) + 10]
    assert oracle.Bitmap.deserialize(synth.from_positions(pos).bytes()).slice() == pos


def test_executor_host_logic_with_oracle_backend():
    ex = Executor(oracle.Oracle(workers=2))
    ex.create_field("f")
    ex.create_field("v", type="int", min=-1000, max=1000)
    assert ex.fields["v"].bsig.bit_depth == 10
    for s in (0, 1):
        ex.load_fragment("f", s, synth.bernoulli(s, 0, 3, 0.05))
        ex.load_fragment("v", s, synth.bsi(s, bit_depth=10, vmin=-1000, vmax=1000, density=0.2))
    a, b = ex.count(Row("f", 0), [0, 1]), ex.count(Row("f", 1), [0, 1])
    i, u, x = (ex.count(op(Row("f", 0), Row("f", 1)), [0, 1]) for op in (Intersect, Union, Xor))
    assert u == a + b - i and x == u - i and ex.count(Difference(Row("f", 0), Row("f", 1)), [0, 1]) == a - i
    s, c = ex.sum("v", [0, 1])
    assert c == ex.count(ex.row_range("v", abi.BSI_NEQ, None), [0, 1])
    lt, ge = ex.count(ex.row_range("v", abi.BSI_LT, 17), [0, 1]), ex.count(ex.row_range("v", abi.BSI_GTE, 17), [0, 1])
    assert lt + ge == c
    assert ex.row_range("v", abi.BSI_GT, 5000).name == "Empty"             # executor.go:1640
    assert ex.row_range("v", abi.BSI_LT, 5000).bsi_op == abi.BSI_NOTNULL    # executor.go:1652
    assert ex.row_range("v", abi.BSI_BETWEEN, -1000, 1000).bsi_op == abi.BSI_NOTNULL  # :1620
    mn, mx = ex.min("v", [0, 1]), ex.max("v", [0, 1])
    assert mn[0] <= mx[0] and mn[1] >= 1 and mx[1] >= 1


_DIST_WORKER = r"""
import os, sys
sys.path.insert(0, {root!r})
import torch.distributed as dist
import oracle
from pilosa_b200 import abi, synth
from pilosa_b200.abi import Intersect, Row, Union
from pilosa_b200.executor import Executor, RowsCall as R

dist.init_process_group("gloo", init_method="tcp://127.0.0.1:{port}", rank=int(sys.argv[1]), world_size=2)
rank = dist.get_rank()
shards = list(range(6))
def build(rank, world):
    ex = Executor(oracle.Oracle(workers=1), rank=rank, world=world)
    ex.create_field("f"); ex.create_field("z"); ex.create_field("v", type="int", min=-500, max=500)
    for s in shards:
        ex.load_fragment("f", s, synth.bernoulli(s, 0, 3, 0.05))
        ex.load_fragment("z", s, synth.zipf(s, n_rows=60, k=2))
        ex.load_fragment("v", s, synth.bsi(s, bit_depth=9, vmin=-500, vmax=500, density=0.01))
    return ex
ex2 = build(rank, 2)            # sharded over two ranks
ex1 = build(0, 1)               # everything on one rank: the expected answers
src = Intersect(Row("f", 0), Row("f", 1))
checks = [
    (ex2.count(src, shards), ex1.count(src, shards)),
    (ex2.count_batch([src, Union(Row("f", 0), Row("f", 2))], shards), ex1.count_batch([src, Union(Row("f", 0), Row("f", 2))], shards)),
    (ex2.sum("v", shards), ex1.sum("v", shards)),
    (ex2.sum("v", shards, Row("f", 1)), ex1.sum("v", shards, Row("f", 1))),
    (ex2.min("v", shards), ex1.min("v", shards)),
    (ex2.max("v", shards), ex1.max("v", shards)),
    (ex2.topn("z", shards, n=5), ex1.topn("z", shards, n=5)),
    (ex2.topn("z", shards, n=5, src=src), ex1.topn("z", shards, n=5, src=src)),
    (ex2.topn("z", shards, n=5, src=src, ids=[1, 2, 3]), ex1.topn("z", shards, n=5, src=src, ids=[1, 2, 3])),
    (ex2.count(ex2.row_range("v", abi.BSI_GT, 0), shards), ex1.count(ex1.row_range("v", abi.BSI_GT, 0), shards)),
    (ex2.count(src, tuple(shards)), ex1.count(src, tuple(shards))),      # tuple shard lists: cached marshalling
    (ex2.count(src, tuple(shards)), ex1.count(src, shards)),
    (ex2.rows("z", shards), ex1.rows("z", shards)),
    (ex2.rows("z", shards, previous=3, limit=4), ex1.rows("z", shards, previous=3, limit=4)),
    (ex2.rows("z", shards, column=(3 << 20) + 5), ex1.rows("z", shards, column=(3 << 20) + 5)),
    (ex2.group_by([R("f"), R("z", limit=5)], shards), ex1.group_by([R("f"), R("z", limit=5)], shards)),
    (ex2.group_by([R("f"), R("f")], shards, filter_call=Row("z", 1), limit=4), ex1.group_by([R("f"), R("f")], shards, filter_call=Row("z", 1), limit=4)),
    (ex2.row_segments(src, shards), [(s_, b_) for s_, b_ in ex1.row_segments(src, shards) if s_ % 2 == rank]),
    # the protobuf Columns field of the rank-local part; a Union wider than one device program through the same merge
    (ex2.row_proto_columns(src, shards), ex1.be.eval_proto_columns(ex1.program(src), [s_ for s_ in shards if s_ % 2 == rank])),
    (ex2.count(Union(*[Row("z", r_) for r_ in range(60)]), shards), ex1.count(Union(*[Row("z", r_) for r_ in range(60)]), shards)),
]
for i, (got, exp) in enumerate(checks):
    assert got == exp, (i, got, exp)
assert sorted(ex2.local_shards(shards)) == [s for s in shards if s % 2 == rank]
dist.barrier()
dist.destroy_process_group()
print("rank", rank, "ok")
"""


def test_two_rank_gloo_shard_partition(tmp_path):
    """world_size-2 gloo run of the shard partition + partial merge: Count/Sum
    all-reduce, Min/Max triple gather, TopN two-pass merge equal the single-rank answers."""
    import socket

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    script = tmp_path / "worker.py"
    script.write_text(_DIST_WORKER.format(root=ROOT, port=port))
    procs = [subprocess.Popen([sys.executable, str(script), str(r)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT) for r in range(2)]
    outs = [p.communicate(timeout=300)[0].decode() for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o
        assert "ok" in o


def test_host_loader_parser_matches_oracle_and_survives_fuzz():
    """The loader's own parser (snapshot + ops-log replay, no GPU) agrees with the oracle's
    restatement of unmarshalPilosaRoaring on valid files and rejects malformed ones cleanly."""
    import random

    from pilosa_b200.engine import EngineError, host_check_roaring

    good = synth.mixed(0, 0, 3).bytes()
    ops = oracle.encode_op(0, [5]) + oracle.encode_op(2, [70000, 1 << 21, (1 << 21) + 9]) + oracle.encode_op(1, [5]) + oracle.encode_op(3, [12])
    for blob in (good, good + ops, synth.bernoulli(2, 0, 2, 0.1).bytes(), synth.bernoulli(2, 0, 2, 1e-4).bytes() + ops):
        ref = oracle.Bitmap.deserialize(blob)
        nc, nb = host_check_roaring(blob)
        assert nb == ref.count() and nc == len(ref.containers())
    for blob, msg in ((b"", "too small"), (b"\x00" * 32, "serialCookie"), (b"\x3c\x30\x01\x00" + b"\x00" * 28, "wrong roaring version"), (good[:40], "truncated"), (good + b"\x09" * 13, "unknown op type"),
                      (good + oracle.encode_op(0, [5])[:-1] + b"\x00", "checksum")):
        with pytest.raises(EngineError, match=msg):
            host_check_roaring(blob)
    rnd = random.Random(2)
    rejected = 0
    for _ in range(400):
        b = bytearray(good if rnd.random() < 0.5 else good[: rnd.randrange(8, len(good))])
        for _ in range(rnd.randrange(1, 8)):
            b[rnd.randrange(min(len(b), 800))] = rnd.randrange(256)
        try:
            host_check_roaring(bytes(b))
        except EngineError:
            rejected += 1
    assert rejected > 200  # and, above all, no crash


def _official_roaring(containers, with_runs=False):
    """Encoder for the OFFICIAL roaring format (the /import-roaring payload): containers = [(key16, kind, payload)]
    with kind 'a' (sorted values), 'b' (1024 words) or 'r' ([(start, last)], needs with_runs)."""
    import struct

    n = len(containers)
    def card(kind, p):
        return len(p) if kind == "a" else sum(bin(w).count("1") for w in p) if kind == "b" else sum(l - s + 1 for s, l in p)
    def body(kind, p):
        if kind == "a":
            return struct.pack(f"<{len(p)}H", *p)
        if kind == "b":
            return struct.pack("<1024Q", *p)
        return struct.pack("<H", len(p)) + b"".join(struct.pack("<HH", s, l - s) for s, l in p)
    hdr = b"".join(struct.pack("<HH", k, card(kind, p) - 1) for k, kind, p in containers)
    bodies = [body(kind, p) for _, kind, p in containers]
    if with_runs:
        bits = bytearray((n + 7) // 8)
        for i, (_, kind, _) in enumerate(containers):
            if kind == "r":
                bits[i // 8] |= 1 << (i % 8)
        return struct.pack("<I", 12347 | ((n - 1) << 16)) + bytes(bits) + hdr + b"".join(bodies)
    out = struct.pack("<II", 12346, n) + hdr
    off = len(out) + 4 * n
    offs = []
    for b in bodies:
        offs.append(off)
        off += len(b)
    return out + struct.pack(f"<{n}I", *offs) + b"".join(bodies)


def test_official_roaring_format_golden_vectors():
    """TestUnmarshalRoaringWithNoErrors / WithErrors (roaring_internal_test.go:3359-3435): the oracle's
    UnmarshalBinary and the loader's host parser read the reference's serialized images alike."""
    import json

    from pilosa_b200.engine import EngineError, host_check_roaring

    g = json.load(open(os.path.join(ROOT, "tests", "golden", "official_roaring.json")))
    for case in g["ok"]:
        blob = bytes.fromhex(case["hex"])
        bm = oracle.Bitmap.deserialize(blob)
        assert bm.count() == case["count"]
        if case["bits"] is not None:
            assert bm.slice() == case["bits"]
        assert host_check_roaring(blob)[1] == case["count"]
    for case in g["errors"]:
        blob = bytes.fromhex(case["hex"])
        if case["error"] is None:  # a Pilosa-format image with no containers: fine
            assert oracle.Bitmap.deserialize(blob).count() == 0 and host_check_roaring(blob) == (0, 0)
            continue
        with pytest.raises(oracle.OracleError) as e1:
            oracle.Bitmap.deserialize(blob)
        assert str(e1.value) == case["error"]
        with pytest.raises(EngineError) as e2:
            host_check_roaring(blob)
        assert case["error"] in str(e2.value)
    # our own encoder against both parsers: arrays + bitmap with offsets; runs (start:length) without
    words = [0] * 1024
    for v in range(0, 65536, 7):
        words[v >> 6] |= 1 << (v & 63)
    a = _official_roaring([(0, "a", [1, 5, 70]), (3, "b", words), (9, "a", list(range(100, 4000, 3)))])
    r = _official_roaring([(0, "r", [(1, 10), (100, 65535)]), (2, "a", [7]), (5, "b", words)], with_runs=True)
    for blob in (a, r):
        bm = oracle.Bitmap.deserialize(blob)
        assert host_check_roaring(blob) == (len(bm.containers()), bm.count())
    assert oracle.Bitmap.deserialize(r).slice()[:12] == list(range(1, 11)) + [100, 101]
    for cut in (9, 20, len(a) - 1, len(a) - 9000):  # truncations are errors in both, never a crash
        with pytest.raises(oracle.OracleError):
            oracle.Bitmap.deserialize(a[:cut])
        with pytest.raises(EngineError):
            host_check_roaring(a[:cut])


def test_plain_c_host_builds_and_fails_loudly_without_a_gpu():
    """examples/count_c.c is a C (not C++) translation unit over include/pilosa_b200.h: it must compile with gcc,
    link against the library alone, and — like every product path — refuse to run without a CUDA device."""
    exe = os.path.join(ROOT, "examples", "count_c")
    assert os.path.exists(exe), "make -C pilosa_b200/csrc builds it"
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu test")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=60)
    assert r.returncode != 0 and "pb200_init" in r.stderr and "C host ok" not in r.stdout


def test_time_quantum_views_golden_table_and_row_time():
    """viewsByTimeRange (time.go:104-181) against the table of TestViewsByTimeRange (time_internal_test.go:87-160),
    and the Row(from=, to=) composition of executor.go:1501-1531 on the host side."""
    import json
    from datetime import datetime

    from pilosa_b200.timeviews import views_by_time, views_by_time_range

    g = json.load(open(os.path.join(ROOT, "tests", "golden", "time_views.json")))
    assert len(g["cases"]) == 13
    ts = lambda s: datetime.strptime(s, "%Y-%m-%d %H:%M")
    for c in g["cases"]:
        assert views_by_time_range("F", ts(c["start"]), ts(c["end"]), c["quantum"]) == c["views"], c["quantum"]
    assert views_by_time("F", ts("2000-01-02 03:00"), "YMDH") == ["F_2000", "F_200001", "F_20000102", "F_2000010203"]  # TestViewsByTime, time_internal_test.go:68-84
    assert views_by_time("F", ts("2000-01-02 03:00"), "D") == ["F_20000102"]
    assert views_by_time_range("F", ts("2001-01-01 00:00"), ts("2001-01-01 00:00"), "YMDH") == []                       # empty range

    ex = Executor(oracle.Oracle(workers=1))
    ex.create_field("t", type="time", time_quantum="YMD")
    ex.create_field("plain")
    for v, seed in (("standard_2001", 1), ("standard_200012", 2), ("standard_20001130", 3), ("standard_2003", 4)):
        ex.load_fragment("t", 0, synth.bernoulli(0, 0, 2, 0.01, seed), view=v)
    full = ex.row_time("t", 1, ts("2000-11-28 00:00"), ts("2003-03-02 00:00"))
    assert full.name == "Union" and [k.view for k in full.children] == ["standard_20001130", "standard_200012", "standard_2001"]
    one = ex.row_time("t", 1, ts("2001-01-01 00:00"), ts("2002-01-01 00:00"))
    assert one.name == "Row" and one.view == "standard_2001"
    assert ex.row_time("t", 1, ts("1990-01-01 00:00"), ts("1991-01-01 00:00")).name == "Empty"   # no such views
    assert ex.row_time("plain", 1, ts("2000-01-01 00:00"), ts("2003-01-01 00:00")).name == "Empty"  # no quantum (:1502-1505)
    parts = [ex.count(abi.Call("Row", field_name="t", row=1, view=k.view), [0]) for k in full.children]
    assert 0 < ex.count(full, [0]) <= sum(parts) and ex.count(full, [0]) >= max(parts)
    assert ex.count(one, [0]) == parts[2]


def test_oracle_proto_row_columns_against_protobuf():
    """The oracle's restatement of internal.Row's marshaller for the Columns field (internal/public.pb.go:745-766):
    a hand-encoded literal and google.protobuf's encoding of the same message (proto3 `repeated uint64 Columns = 1`,
    packed) for column lists that cross every varint-length step below 2^29."""
    from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

    from pilosa_b200.abi import Row
    from pilosa_b200.executor import Executor

    fdp = descriptor_pb2.FileDescriptorProto(name="public_min.proto", package="internal", syntax="proto3")
    m = fdp.message_type.add(name="Row")
    m.field.add(name="Columns", number=1, type=descriptor_pb2.FieldDescriptorProto.TYPE_UINT64, label=descriptor_pb2.FieldDescriptorProto.LABEL_REPEATED)
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fdp)
    RowMsg = message_factory.GetMessageClass(pool.FindMessageTypeByName("internal.Row"))

    orc = oracle.Oracle(workers=1)
    c = Executor(orc)
    c.create_field("f")
    fam = c.family("f")
    for col in (1, 300, 1 << 20):
        orc.set_bit(fam, col >> 20, 7, col)
    # 1 -> 01; 300 -> AC 02; 2^20 -> 80 80 40; tag 0A, length 6
    assert c.row_proto_columns(Row("f", 7), [0, 1]) == bytes([0x0A, 0x06, 0x01, 0xAC, 0x02, 0x80, 0x80, 0x40])
    assert c.row_proto_columns(Row("f", 7), [1]) == bytes([0x0A, 0x03, 0x80, 0x80, 0x40])
    assert c.row_proto_columns(Row("f", 8), [0, 1]) == b""  # len(m.Columns) == 0: nothing is written (:750)
    cols = [0, 1, 127, 128, 16383, 16384, (1 << 20) - 1, 1 << 20, (1 << 21) - 1, 1 << 21, (255 << 20) + 5, (1 << 28) - 1, 1 << 28, (300 << 20) + 12345]
    for col in cols:
        orc.set_bit(fam, col >> 20, 9, col)
    shards = sorted({col >> 20 for col in cols})
    assert c.row_proto_columns(Row("f", 9), shards) == RowMsg(Columns=cols).SerializeToString()
    many = list(range(0, 70000, 7))
    for col in many:
        orc.set_bit(fam, 0, 10, col)
    assert c.row_proto_columns(Row("f", 10), [0]) == RowMsg(Columns=many).SerializeToString()  # 2-byte length prefix and beyond


def test_min_max_views_and_time_of_view():
    """TestMinMaxViews (time_internal_test.go:168-221) and TestTimeOfView (:223-293): the tables, literally."""
    from datetime import datetime

    from pilosa_b200.timeviews import min_max_views, time_of_view

    for views, q, mn, mx in (([""], "Y", "", ""),
                             (["std_2019", "std_2020", "std_202002", "std_202002", "std_2022"], "Y", "std_2019", "std_2022"),
                             (["std_201902", "std_201901"], "M", "std_201901", "std_201902"),
                             (["std_201902", "std_201901"], "D", "", ""),
                             (["std_20190201"], "D", "std_20190201", "std_20190201"),
                             (["foo", "bar"], "D", "", "")):
        assert min_max_views(views, q) == (mn, mx), (views, q)
    for view, exp, exp_adj in (("std_2019", datetime(2019, 1, 1), datetime(2020, 1, 1)),
                               ("std_201902", datetime(2019, 2, 1), datetime(2019, 3, 1)),
                               ("std_20190203", datetime(2019, 2, 3), datetime(2019, 2, 4)),
                               ("std_2019020308", datetime(2019, 2, 3, 8), datetime(2019, 2, 3, 9))):
        assert time_of_view(view, False) == exp and time_of_view(view, True) == exp_adj, view
    for view in ("foo", "std_201902030801"):
        for adj in (False, True):
            with pytest.raises(ValueError, match=f"invalid time format on view: {view}"):
                time_of_view(view, adj)
    assert time_of_view("", False) is None  # the zero time (:280-282)


def test_oracle_worker_pool_batch_matches_single_calls():
    """The oracle's persistent worker pool (executor.go:83-107): a batch of Count calls pushed as queries x shards jobs
    gives the per-call results, in both modes, for any pool size, and the pool is reused across calls."""
    import ctypes as C

    import oracle
    from pilosa_b200 import abi, synth
    from pilosa_b200.abi import Intersect, Row, Union
    from pilosa_b200.executor import Executor

    orc = oracle.Oracle(workers=4)
    ex = Executor(orc)
    ex.create_field("f")
    ex.create_field("m")
    shards = [0, 1, 2, 3, 4]
    for s in shards:
        ex.load_fragment("f", s, synth.bernoulli(s, 0, 6, 0.10))
        ex.load_fragment("m", s, synth.mixed(s, 0, 4))
    calls = [Intersect(Row("f", a), Row("f", b)) for a in range(6) for b in range(a + 1, 6)] + [Union(Row("m", 0), Row("m", 1), Row("m", 3)), Row("m", 2)]
    progs = [ex.program(c) for c in calls]
    flat = (abi.Node * sum(len(p) for p in progs))()
    off = (C.c_uint32 * (len(progs) + 1))()
    k = 0
    for i, p in enumerate(progs):
        off[i] = k
        for nd in p:
            flat[k] = nd
            k += 1
    off[len(progs)] = k
    sh = abi.u64_array(shards)
    orc.set_workers(1)
    want = [orc.count(p, shards) for p in progs]
    for workers in (1, 3, 8):
        orc.set_workers(workers)
        for fused in (False, True):
            for _ in range(3):
                got = orc.count_batch_raw(flat, off, len(progs), sh, len(shards), fused)
                assert list(got[: len(progs)]) == want
        assert [orc.count(p, shards, True) for p in progs[:4]] == want[:4]
