#!/usr/bin/env python
"""bench_configs.py — the other BASELINE.json configs (C2..C5 of SURVEY.md §8d) on one GPU.

bench.py carries the headline metric (C1 batched). This script measures, for each remaining
config, the engine through the C ABI (kernel time from the library's CUDA events, wall time of
the whole call, algorithmic GB/s) next to the CPU oracle (reference restatement, all host
cores) on a bounded sample of the same shards, and checks the sampled answers bit-exactly.
One JSON line per query, plus a markdown table on stderr for BASELINE.md.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import time
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from pilosa_b200 import abi, synth  # noqa: E402
from pilosa_b200.abi import Difference, Intersect, Row, Union, Xor  # noqa: E402
from pilosa_b200.executor import Executor  # noqa: E402

SEED = 20190416


def load_parallel(gen, shards, sinks, workers):
    """gen(shard) -> FragBytes, loaded into every sink(shard, fb)."""
    with ThreadPoolExecutor(max_workers=workers) as tp:
        for s, fb in zip(shards, tp.map(gen, shards)):
            for sink in sinks(s):
                sink(s, fb)


def timed(eng, fn, reps):
    fn()  # warm-up (also grows scratch buffers)
    fn()
    ks, ws, res = [], [], None
    for _ in range(reps):
        t0 = time.perf_counter()
        res = fn()
        ws.append((time.perf_counter() - t0) * 1e3)
        ks.append(eng.last_timing().kernel_ms)
    tm = eng.last_timing()
    return res, statistics.median(ks), statistics.median(ws), int(tm.algo_bytes), int(tm.launches)


def cpu_timed(fn, min_seconds=3.0):
    fn()
    t0 = time.perf_counter()
    n = 0
    res = None
    while n < 1 or time.perf_counter() - t0 < min_seconds:
        res = fn()
        n += 1
    return res, (time.perf_counter() - t0) / n * 1e3


def emit(rows, name, query, n_shards, kernel_ms, wall_ms, algo_bytes, launches, cpu_ms, cpu_shards, ok, extra=None):
    gbs = algo_bytes / (kernel_ms * 1e-3) / 1e9 if kernel_ms > 0 else 0.0
    cpu_full_ms = cpu_ms * n_shards / max(1, cpu_shards)
    line = {"config": name, "query": query, "shards": n_shards, "kernel_ms": round(kernel_ms, 4), "call_ms": round(wall_ms, 4),
            "algo_bytes": algo_bytes, "algo_GBps": round(gbs, 1), "launches": launches,
            "cpu_ms_sample": round(cpu_ms, 3), "cpu_sample_shards": cpu_shards, "cpu_ms_scaled_to_all_shards": round(cpu_full_ms, 2),
            "cpu_cores": os.cpu_count(), "speedup_call_vs_cpu": round(cpu_full_ms / wall_ms, 1) if wall_ms > 0 else None,
            "parity": "ok" if ok else "MISMATCH"}
    if extra:
        line.update(extra)
    print(json.dumps(line), flush=True)
    rows.append(line)
    if not ok:
        raise SystemExit(f"parity mismatch in {name}: {query}")


def prop(rows, config, name, ok, detail=""):
    """A size-independent identity checked on the GPU results at the config's FULL size (the oracle only
    covers a shard sample there)."""
    line = {"config": config, "property": name, "ok": bool(ok), "detail": detail}
    print(json.dumps(line), flush=True)
    if not ok:
        raise SystemExit(f"property violated in {config}: {name} {detail}")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="c1,c2,c3,c4,c5,gb")
    ap.add_argument("--gb-shards", type=int, default=256)
    ap.add_argument("--reps", type=int, default=7)
    ap.add_argument("--c2-shards", type=int, default=1024)
    ap.add_argument("--c3-shards", type=int, default=1024)
    ap.add_argument("--c3-rows", type=int, default=10000)
    ap.add_argument("--c4-shards", type=int, default=2048)
    ap.add_argument("--c5-shards", default="64,512,8192")
    ap.add_argument("--cpu-shards", type=int, default=32)
    args = ap.parse_args()

    import oracle
    from pilosa_b200.engine import Engine

    cores = os.cpu_count() or 8
    eng = Engine(0)
    g = Executor(eng)
    c = Executor(oracle.Oracle(workers=cores))
    rows = []
    cfgs = args.configs.split(",")

    def both(name, **kw):
        for ex in (g, c):
            ex.create_field(name, **kw)

    def sinks_for(field, cpu_set):
        def sinks(s):
            out = [lambda s_, fb: g.load_fragment(field, s_, fb)]
            if s in cpu_set:
                out.append(lambda s_, fb: c.load_fragment(field, s_, fb))
            return out
        return sinks

    # ---------------- C1 (single query): latency of ONE Count(Intersect(Row(f=1),Row(f=2))), 16 shards ----------------
    # Launch/latency-bound: 4 MiB of operands stay L2-resident when the same pair is repeated, so
    # the GB/s figure here is an L2 number, labelled as such; the HBM-bound batched figure is bench.py's.
    if "c1" in cfgs:
        shards = tuple(range(16))
        both("f")
        load_parallel(lambda s: synth.bernoulli(s, 0, 8, 0.10, SEED), shards, sinks_for("f", set(shards)), cores)
        call = Intersect(Row("f", 1), Row("f", 2))
        _, k, wms, ab, ln = timed(eng, lambda: g.count(call, shards), 51)
        got = g.count(call, shards)
        exp, cms = cpu_timed(lambda: c.count(call, shards), 2.0)
        emit(rows, "C1", "single Count(Intersect(Row(f=1),Row(f=2))), same pair repeated (L2-resident)", 16, k, wms, ab, ln, cms, 16, got == exp)
        prog = g.program(call)
        _, k, wms, ab, ln = timed(eng, lambda: eng.count_batch([prog] * 1024, shards), 11)
        emit(rows, "C1", "batch of 1024 x the same pair (L2-resident; per-query figures)", 16, k / 1024, wms / 1024, ab // 1024, ln, cms, 16, True)

    # ---------------- C2: Count(Union/Difference/Xor) over 4 rows, mixed containers ----------------
    if "c2" in cfgs:
        shards = tuple(range(args.c2_shards))
        cpu_sh = shards[: args.cpu_shards]
        both("m")
        t0 = time.perf_counter()
        load_parallel(lambda s: synth.mixed(s, 0, 4, SEED), shards, sinks_for("m", set(cpu_sh)), cores)
        print(f"# c2 loaded in {time.perf_counter() - t0:.1f}s", file=sys.stderr)
        leaves = [Row("m", r) for r in range(4)]
        for qname, call in (("Count(Union(r0..r3))", Union(*leaves)), ("Count(Difference(r0,r1,r2,r3))", Difference(*leaves)),
                            ("Count(Xor(r0..r3))", Xor(*leaves)), ("Count(Intersect(r0,r1))", Intersect(leaves[0], leaves[1]))):
            _, k, wms, ab, ln = timed(eng, lambda: g.count(call, shards), args.reps)
            got = g.count(call, cpu_sh)
            exp, cms = cpu_timed(lambda: c.count(call, cpu_sh))
            emit(rows, "C2", qname, len(shards), k, wms, ab, ln, cms, len(cpu_sh), got == exp)

        # the same queries with PB200_OPT_DENSIFY_MIN_BYTES = 1024 (second context: large arrays / runs kept as bitmaps)
        eng_d = Engine(0)
        eng_d.set_option(Engine.OPT_DENSIFY_MIN_BYTES, 1024)
        gd = Executor(eng_d)
        gd.create_field("m")
        load_parallel(lambda s: synth.mixed(s, 0, 4, SEED), shards, lambda s: [lambda s_, fb: gd.load_fragment("m", s_, fb)], cores)
        hbm = [sum(e.frag_stat(x.family("m"), s)[2] for s in shards) for e, x in ((eng, g), (eng_d, gd))]
        for qname, call in (("Count(Union(r0..r3))", Union(*leaves)), ("Count(Difference(r0,r1,r2,r3))", Difference(*leaves)),
                            ("Count(Xor(r0..r3))", Xor(*leaves)), ("Count(Intersect(r0,r1))", Intersect(leaves[0], leaves[1]))):
            _, k, wms, ab, ln = timed(eng_d, lambda: gd.count(call, shards), args.reps)
            same = gd.count(call, shards) == g.count(call, shards)
            exp, cms = cpu_timed(lambda: c.count(call, cpu_sh), 1.0)
            emit(rows, "C2", qname + " [densify >= 1024 B]", len(shards), k, wms, ab, ln, cms, len(cpu_sh), same and gd.count(call, cpu_sh) == exp,
                 {"hbm_payload_bytes": hbm[1], "hbm_payload_bytes_as_loaded": hbm[0]})
        eng_d.close()
        if "c2" in cfgs:
            a, b = leaves[0], leaves[1]
            ca, cb = g.count(a, shards), g.count(b, shards)
            ci, cu, cx, cd = (g.count(op(a, b), shards) for op in (Intersect, Union, Xor, Difference))
            prop(rows, "C2", "|a|+|b| = |a or b| + |a and b|", ca + cb == cu + ci, f"{ca}+{cb} vs {cu}+{ci}")
            prop(rows, "C2", "|a xor b| = |a or b| - |a and b|", cx == cu - ci)
            prop(rows, "C2", "|a - b| = |a| - |a and b|", cd == ca - ci)
            per = eng.count_shards(g.program(Union(*leaves)), shards)
            prop(rows, "C2", "sum of per-shard counts = count over all shards", sum(per) == g.count(Union(*leaves), shards))
            u4, x4 = g.count(Union(*leaves), shards), g.count(Xor(*leaves), shards)
            prop(rows, "C2", "Xor(r0..r3) <= Union(r0..r3) <= sum of rows", x4 <= u4 <= sum(g.count(l, shards) for l in leaves))

    # ---------------- C3: TopN(f, n=50) over 10k rows with an Intersect filter ----------------
    if "c3" in cfgs:
        shards = tuple(range(args.c3_shards))
        cpu_sh = shards[: max(4, args.cpu_shards // 4)]
        both("z")
        both("a")
        t0 = time.perf_counter()
        load_parallel(lambda s: synth.zipf(s, args.c3_rows, 8, 1.6, 50.0, SEED), shards, sinks_for("z", set(cpu_sh)), cores)
        load_parallel(lambda s: synth.bernoulli(s, 1, 2, 0.10, SEED), shards, sinks_for("a", set(cpu_sh)), cores)
        print(f"# c3 loaded in {time.perf_counter() - t0:.1f}s", file=sys.stderr)
        src = Intersect(Row("a", 1), Row("a", 2))
        # SURVEY 8d asks for both byte figures: A_pruned = what fragment.top's walk visits (the engine's default: it prunes
        # by the reference's own rule on the device), A_full = every ranked row (pruning switched off)
        for qname, kw, prune in ((f"TopN(z, Intersect(Row(a=1),Row(a=2)), n=50) [{args.c3_rows} rows] A_pruned", dict(n=50, src=src), 1),
                                 (f"TopN(z, Intersect(Row(a=1),Row(a=2)), n=50) [{args.c3_rows} rows] A_full (pruning off)", dict(n=50, src=src), 0),
                                 ("TopN(z, n=50) [no source: rank cache only]", dict(n=50), 1)):
            eng.set_option(3, prune)  # PB200_OPT_TOPN_PRUNE
            _, k, wms, ab, ln = timed(eng, lambda: g.topn("z", shards, **kw), max(3, args.reps // 2))
            got = g.topn("z", cpu_sh, **kw)
            eng.set_option(3, 1)
            exp, cms = cpu_timed(lambda: c.topn("z", cpu_sh, **kw))
            emit(rows, "C3", qname, len(shards), k, wms, ab, ln, cms, len(cpu_sh), got == exp)

        # n-ary operations wider than one device program (32 leaves): a Union of rows is one accumulating pass (K13)
        for lo, hi in ((0, 100), (1000, 1400)):
            call = Union(*[Row("z", r) for r in range(lo, hi)])
            _, k, wms, ab, ln = timed(eng, lambda: g.count(call, shards), max(3, args.reps // 2))
            got = g.count(call, cpu_sh)
            exp, cms = cpu_timed(lambda: c.count(call, cpu_sh))
            emit(rows, "C3", f"Count(Union(Row(z={lo}) .. Row(z={hi - 1}))) [{hi - lo} operands]", len(shards), k, wms, ab, ln, cms, len(cpu_sh), got == exp)
        # ... Xor / Intersect of more than 32 rows: chunks of 32 folded into dense temp rows by the fold evaluator, then combined
        for opname, op, lo, hi in (("Xor", Xor, 0, 100), ("Intersect", Intersect, 0, 100), ("Xor", Xor, 1000, 1400)):
            call = op(*[Row("z", r) for r in range(lo, hi)])
            _, k, wms, ab, ln = timed(eng, lambda: g.count(call, shards), max(3, args.reps // 2))
            got = g.count(call, cpu_sh)
            exp, cms = cpu_timed(lambda: c.count(call, cpu_sh))
            emit(rows, "C3", f"Count({opname}(Row(z={lo}) .. Row(z={hi - 1}))) [{hi - lo} operands]", len(shards), k, wms, ab, ln, cms, len(cpu_sh), got == exp)
        wide = Union(*[Row("z", r) for r in range(0, 64)])
        prop(rows, "C3", "Count(Union of 64 rows) = sum over inclusion-exclusion halves: |A u B| = |A| + |B| - |A n B|",
             g.count(wide, shards) == g.count(Union(*[Row("z", r) for r in range(32)]), shards) + g.count(Union(*[Row("z", r) for r in range(32, 64)]), shards)
             - g.count(Intersect(Union(*[Row("z", r) for r in range(32)]), Union(*[Row("z", r) for r in range(32, 64)])), shards))

        if "c3" in cfgs:
            top = g.topn("z", shards, n=50, src=src)
            exact = g.count_batch([Intersect(src, Row("z", i)) for i, _ in top], shards)
            prop(rows, "C3", "TopN counts = Count(Intersect(src, Row)) of the returned ids", [c_ for _, c_ in top] == exact)
            prop(rows, "C3", "TopN result sorted by count desc", all(top[i][1] >= top[i + 1][1] for i in range(len(top) - 1)))
            again = g.topn("z", shards, n=50, src=src, ids=[i for i, _ in top])
            prop(rows, "C3", "TopN(ids = its own result) returns the same pairs", sorted(again) == sorted(top))

    # ---------------- C4: BSI int32: Sum / Min / Max / Range ----------------
    if "c4" in cfgs:
        shards = tuple(range(args.c4_shards))
        cpu_sh = shards[: max(4, args.cpu_shards // 4)]
        both("v", type="int", min=-(1 << 31), max=(1 << 31) - 1)
        both("fl")
        t0 = time.perf_counter()
        load_parallel(lambda s: synth.bsi(s, 32, -(1 << 31), (1 << 31) - 1, 1.0, SEED), shards, sinks_for("v", set(cpu_sh)), cores)
        load_parallel(lambda s: synth.bernoulli(s, 1, 1, 0.10, SEED), shards, sinks_for("fl", set(cpu_sh)), cores)
        print(f"# c4 loaded in {time.perf_counter() - t0:.1f}s", file=sys.stderr)
        qs = [("Sum(field=v)", lambda ex, sh: ex.sum("v", sh)), ("Sum(Row(fl=1), field=v)", lambda ex, sh: ex.sum("v", sh, Row("fl", 1))),
              ("Min(field=v)", lambda ex, sh: ex.min("v", sh)), ("Max(field=v)", lambda ex, sh: ex.max("v", sh))]
        for x in (0, 1 << 30, -(1 << 30), 12345):
            qs.append((f"Count(Row(v > {x}))", lambda ex, sh, x=x: ex.count(ex.row_range("v", abi.BSI_GT, x), sh)))
        qs.append(("Count(Row(-1000 < v < 1000))", lambda ex, sh: ex.count(ex.row_range("v", abi.BSI_BETWEEN, -999, 999), sh)))
        for qname, fn in qs:
            _, k, wms, ab, ln = timed(eng, lambda: fn(g, shards), args.reps)
            got = fn(g, cpu_sh)
            exp, cms = cpu_timed(lambda: fn(c, cpu_sh))
            emit(rows, "C4", qname, len(shards), k, wms, ab, ln, cms, len(cpu_sh), got == exp)
        nn = g.count(g.row_range("v", abi.BSI_NEQ, None), shards)  # not null
        for x in (0, 12345, -(1 << 30)):
            gt, le = g.count(g.row_range("v", abi.BSI_GT, x), shards), g.count(g.row_range("v", abi.BSI_LTE, x), shards)
            eq, ne = g.count(g.row_range("v", abi.BSI_EQ, x), shards), g.count(g.row_range("v", abi.BSI_NEQ, x), shards)
            prop(rows, "C4", f"count(v > {x}) + count(v <= {x}) = count(v != null)", gt + le == nn, f"{gt}+{le} vs {nn}")
            prop(rows, "C4", f"count(v == {x}) + count(v != {x}) = count(v != null)", eq + ne == nn)
        (mn, _), (mx, _) = g.min("v", shards), g.max("v", shards)
        prop(rows, "C4", "count(v < Min) = 0 and count(v > Max) = 0",
             g.count(g.row_range("v", abi.BSI_LT, mn), shards) == 0 and g.count(g.row_range("v", abi.BSI_GT, mx), shards) == 0)
        s_all, c_all = g.sum("v", shards)
        s_f, c_f = g.be.bsi_sum(g.family("v", "bsig_v"), 32, g.program(Row("fl", 1)), shards, flags=1)        # strict: filter applies to negatives too
        s_n, c_n = g.be.bsi_sum(g.family("v", "bsig_v"), 32, g.program(Difference(g.row_range("v", abi.BSI_NEQ, None), Row("fl", 1))), shards, flags=1)
        prop(rows, "C4", "Sum(filter) + Sum(not filter) = Sum(all) (strict sums), counts too", s_f + s_n == s_all and c_f + c_n == c_all == nn)
        bw = g.count(g.row_range("v", abi.BSI_BETWEEN, -999, 999), shards)
        prop(rows, "C4", "count(-1000 < v < 1000) = count(v < 1000) - count(v <= -1000)",
             bw == g.count(g.row_range("v", abi.BSI_LT, 1000), shards) - g.count(g.row_range("v", abi.BSI_LTE, -1000), shards))
        # materialised Row(v > x): every shard's result typed + serialised on the device (K11), packed streams copied
        # to the host (the D2H of ~128 KiB per shard dominates the call; kernel ms is evaluation + serialisation)
        for x in (0, 12345, (1 << 31) - 4096):
            call = g.row_range("v", abi.BSI_GT, x)
            prog = g.program(call)
            res, k, wms, ab, ln = timed(eng, lambda: eng.eval_shards_raw(prog, shards), max(3, args.reps // 2))
            out_bytes = int(res[1][len(shards)])
            got = g.row_segments(call, cpu_sh)
            exp, cms = cpu_timed(lambda: c.row_segments(call, cpu_sh))
            emit(rows, "C4", f"Row(v > {x}) materialised", len(shards), k, wms, ab, ln, cms, len(cpu_sh), got == exp, {"result_bytes": out_bytes})

        # the same kind of result as the protobuf the client answer carries: Columns of internal.Row, sized and
        # varint-encoded on the device (K12), one message over all shards
        for label, call in ((f"Row(v > {(1 << 31) - 4096})", g.row_range("v", abi.BSI_GT, (1 << 31) - 4096)),
                            (f"Row(v > {(1 << 31) - (1 << 24)})", g.row_range("v", abi.BSI_GT, (1 << 31) - (1 << 24))),
                            ("Row(-1000 < v < 1000)", g.row_range("v", abi.BSI_BETWEEN, -999, 999))):
            prog, sh_arr = g.program(call), abi.u64_array(shards)
            res, k, wms, ab, ln = timed(eng, lambda: eng.eval_proto_columns_raw(prog, sh_arr), max(3, args.reps // 2))
            got = g.row_proto_columns(call, cpu_sh)
            exp, cms = cpu_timed(lambda: c.row_proto_columns(call, cpu_sh))
            emit(rows, "C4", f"{label} as protobuf Row.Columns", len(shards), k, wms, ab, ln, cms, len(cpu_sh), got == exp, {"result_bytes": int(res[1])})

    # ---------------- C5: sparse Count(Intersect) sweep ----------------
    if "c5" in cfgs:
        sweep = [int(x) for x in args.c5_shards.split(",")]
        nmax = max(sweep)
        shards = tuple(range(nmax))
        cpu_sh = shards[: min(nmax, 512)]
        both("sp")
        t0 = time.perf_counter()
        load_parallel(lambda s: synth.bernoulli(s, 1, 2, 1e-4, SEED), shards, sinks_for("sp", set(cpu_sh)), cores)
        print(f"# c5 loaded in {time.perf_counter() - t0:.1f}s", file=sys.stderr)
        call = Intersect(Row("sp", 1), Row("sp", 2))
        for n in sweep:
            sh = shards[:n]
            prog, sh_arr = g.program(call), abi.u64_array(sh)  # marshalled once: call ms is the C-ABI call, not list conversion
            _, k, wms, ab, ln = timed(eng, lambda: eng.count(prog, sh_arr), args.reps)
            csh = sh[: len(cpu_sh)]
            got = g.count(call, csh)
            exp, cms = cpu_timed(lambda: c.count(call, csh), 1.0)
            emit(rows, "C5", f"Count(Intersect(Row(sp=1),Row(sp=2))) p=1e-4, {n} shards", n, k, wms, ab, ln, cms, len(csh), got == exp)
            # batched: 256 copies of the query in one call (what a busy server would issue)
            flat, off = eng.pack_programs([prog] * 256)
            _, k, wms, ab, ln = timed(eng, lambda: eng.count_batch_raw(flat, off, 256, sh_arr, n), args.reps)
            emit(rows, "C5", f"batch of 256 x same query, {n} shards (per-query figures)", n, k / 256, wms / 256, ab // 256, ln, cms, len(csh), True)

        prog, sh_arr = g.program(Row("sp", 1)), abi.u64_array(shards)
        res, k, wms, ab, ln = timed(eng, lambda: eng.eval_proto_columns_raw(prog, sh_arr), max(3, args.reps // 2))
        got = g.row_proto_columns(Row("sp", 1), cpu_sh)
        exp, cms = cpu_timed(lambda: c.row_proto_columns(Row("sp", 1), cpu_sh))
        emit(rows, "C5", f"Row(sp=1) p=1e-4 as protobuf Row.Columns, {nmax} shards", nmax, k, wms, ab, ln, cms, len(cpu_sh), got == exp, {"result_bytes": int(res[1])})

    # ---------------- GB: GroupBy over two / three set fields ----------------
    if "gb" in cfgs:
        from pilosa_b200.executor import RowsCall as R
        shards = tuple(range(args.gb_shards))
        cpu_sh = shards[: max(4, args.cpu_shards // 4)]
        for name, nrows, p_ in (("ga", 16, 0.10), ("gb", 64, 0.02), ("gc", 8, 0.25)):
            both(name)
            load_parallel(lambda s, nrows=nrows, p_=p_, name=name: synth.bernoulli(s, 0, nrows, p_, SEED + len(name) + nrows), shards,
                          sinks_for(name, set(cpu_sh)), cores)
        both("gf")
        load_parallel(lambda s: synth.bernoulli(s, 1, 1, 0.5, SEED + 99), shards, sinks_for("gf", set(cpu_sh)), cores)
        for qname, ch, kw in (("GroupBy(Rows(ga), Rows(gb)) [16 x 64 groups]", [R("ga"), R("gb")], {}),
                              ("GroupBy(Rows(ga), Rows(gb), filter=Row(gf=1))", [R("ga"), R("gb")], dict(filter_call=Row("gf", 1))),
                              ("GroupBy(Rows(gc), Rows(ga), Rows(gb)) [8 x 16 x 64 groups]", [R("gc"), R("ga"), R("gb")], {})):
            _, k, wms, ab, ln = timed(eng, lambda: g.group_by(ch, shards, **kw), max(3, args.reps // 2))
            got = g.group_by(ch, cpu_sh, **kw)
            exp, cms = cpu_timed(lambda: c.group_by(ch, cpu_sh, **kw))
            emit(rows, "GB", qname, len(shards), k, wms, ab, ln, cms, len(cpu_sh), got == exp, {"groups": len(got)})

    print("\n| config | query | shards | kernel ms | call ms | algorithmic GB/s | CPU ms (all shards, %d cores) | call speed-up |" % cores, file=sys.stderr)
    print("|---|---|---|---|---|---|---|---|", file=sys.stderr)
    for r in rows:
        print(f"| {r['config']} | `{r['query']}` | {r['shards']} | {r['kernel_ms']} | {r['call_ms']} | {r['algo_GBps']} | "
              f"{r['cpu_ms_scaled_to_all_shards']} | {r['speedup_call_vs_cpu']}x |", file=sys.stderr)
    eng.close()


if __name__ == "__main__":
    main()
