#!/usr/bin/env python
"""bench_configs_dist.py — BASELINE.json configs C2..C5 over N GPUs of one box, through the C ABI only.

    python bench_configs_dist.py --gpus N          (ONE process; N = 1, 2, 4, 8)

Every query is ONE call on the pb200_multi handle (include/pilosa_b200.h): the library partitions the caller's shard
list (shard s on device s % N), runs the owners concurrently — or, for a small call, one device over its peers' HBM —
and returns the reduced answer, which is what executor.mapReduce does in the reference (executor.go:2458-2611). No
torch.distributed, no NCCL and no Python merge is involved. Per query one JSON line: wall ms of the call (median),
the slowest device's kernel ms, and the RESULT, so that runs at different N can be compared for equality — the
full-size parity check across GPU counts.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import statistics
import sys
import time
import zlib
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from pilosa_b200 import abi, synth  # noqa: E402
from pilosa_b200.abi import Difference, Intersect, Row, Union, Xor  # noqa: E402
from pilosa_b200.engine import MultiEngine, device_count  # noqa: E402
from pilosa_b200.executor import Executor  # noqa: E402

SEED = 20190416


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=0, help="devices 0..N-1 (default: all visible)")
    ap.add_argument("--configs", default="c2,c3,c4,c5")
    ap.add_argument("--reps", type=int, default=9)
    ap.add_argument("--c2-shards", type=int, default=1024)
    ap.add_argument("--c3-shards", type=int, default=1024)
    ap.add_argument("--c3-rows", type=int, default=10000)
    ap.add_argument("--c4-shards", type=int, default=2048)
    ap.add_argument("--c5-shards", type=int, default=8192)
    args = ap.parse_args()
    n = args.gpus or device_count()
    eng = MultiEngine(list(range(n)))
    g = Executor(eng)
    cores = max(1, min(16, os.cpu_count() or 8))
    cfgs = args.configs.split(",")

    def load(field, gen, shards):
        with ThreadPoolExecutor(max_workers=cores) as tp:
            for s, fb in zip(shards, tp.map(gen, shards)):
                g.load_fragment(field, s, fb)

    def run(config, name, fn, shards, reps=None, show=lambda r: r):
        reps = reps or args.reps
        fn(); fn()
        walls, kern = [], []
        for _ in range(reps):
            t0 = time.perf_counter()
            res = fn()
            walls.append((time.perf_counter() - t0) * 1e3)
            kern.append(eng.last_timing().kernel_ms)
        print(json.dumps({"config": config, "query": name, "n_gpus": n, "shards": len(shards), "call_ms": round(statistics.median(walls), 4),
                          "call_ms_min": round(min(walls), 4), "call_ms_max": round(max(walls), 4), "kernel_ms_slowest_device": round(statistics.median(kern), 4), "result": show(res)}), flush=True)

    def counter(call, shards):
        """One raw pb200_multi_count call with everything marshalled beforehand."""
        prog = g.program(call)
        sh = abi.u64_array(shards)
        out = C.c_uint64()
        lib, h, np_, ns = eng.lib.pb200_count, eng._h, len(prog), len(shards)

        def fn():
            rc = lib(h, prog, np_, sh, ns, C.byref(out))
            if rc:
                eng._check(rc)
            return out.value
        return fn

    if "c2" in cfgs:
        shards = tuple(range(args.c2_shards))
        g.create_field("m")
        load("m", lambda s: synth.mixed(s, 0, 4, SEED), shards)
        leaves = [Row("m", r) for r in range(4)]
        for qname, call in (("Count(Union(r0..r3))", Union(*leaves)), ("Count(Difference(r0,r1,r2,r3))", Difference(*leaves)),
                            ("Count(Xor(r0..r3))", Xor(*leaves))):
            run("C2", qname, counter(call, shards), shards)
    if "c3" in cfgs:
        shards = tuple(range(args.c3_shards))
        g.create_field("z")
        g.create_field("a")
        load("z", lambda s: synth.zipf(s, args.c3_rows, 8, 1.6, 50.0, SEED), shards)
        load("a", lambda s: synth.bernoulli(s, 1, 2, 0.10, SEED), shards)
        src = Intersect(Row("a", 1), Row("a", 2))
        run("C3", f"TopN(z, Intersect(Row(a=1),Row(a=2)), n=50) [{args.c3_rows} rows]", lambda: g.topn("z", shards, n=50, src=src), shards, max(3, args.reps // 2))
    if "c4" in cfgs:
        shards = tuple(range(args.c4_shards))
        g.create_field("v", type="int", min=-(1 << 31), max=(1 << 31) - 1)
        g.create_field("fl")
        load("v", lambda s: synth.bsi(s, 32, -(1 << 31), (1 << 31) - 1, 1.0, SEED), shards)
        load("fl", lambda s: synth.bernoulli(s, 1, 1, 0.10, SEED), shards)
        run("C4", "Sum(field=v)", lambda: g.sum("v", shards), shards)
        run("C4", "Sum(Row(fl=1), field=v)", lambda: g.sum("v", shards, Row("fl", 1)), shards)
        run("C4", "Min(field=v)", lambda: g.min("v", shards), shards)
        run("C4", "Max(field=v)", lambda: g.max("v", shards), shards)
        for x in (0, 12345):
            run("C4", f"Count(Row(v > {x}))", counter(g.row_range("v", abi.BSI_GT, x), shards), shards)
    if "c5" in cfgs:
        shards = tuple(range(args.c5_shards))
        g.create_field("sp")
        load("sp", lambda s: synth.bernoulli(s, 1, 2, 1e-4, SEED), shards)
        call = Intersect(Row("sp", 1), Row("sp", 2))
        sizes = sorted({k for k in (64, 512, args.c5_shards) if k <= args.c5_shards})
        for k in sizes:
            sh = shards[:k]
            run("C5", f"Count(Intersect(Row(sp=1),Row(sp=2))) p=1e-4, {k} shards", counter(call, sh), sh)
            flat, off = eng.pack_programs([g.program(call)] * 256)   # marshalled once: the call is the C ABI call
            sh_arr = abi.u64_array(sh)
            run("C5", f"batch of 256 x same query, {k} shards (whole batch)", lambda: list(eng.count_batch_raw(flat, off, 256, sh_arr, len(sh))[:2]), sh)
        # a materialised sparse row over all shards: owners evaluate, the first device gathers and emits the protobuf Columns field
        prog_row, sh_all = g.program(Row("sp", 1)), abi.u64_array(shards)   # marshalled once; the message lands in a reused page-locked buffer
        run("C5", f"Row(sp=1) as protobuf Columns, {len(shards)} shards", lambda: eng.eval_proto_columns_raw(prog_row, sh_all), shards, 5,
            show=lambda r: {"bytes": r[1], "crc32": zlib.crc32(C.string_at(r[0], r[1]))})
    eng.close()


if __name__ == "__main__":
    main()
