#!/usr/bin/env python
"""bench_configs_dist.py — BASELINE.json configs C2..C5 with the shards partitioned over N GPUs of one box.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench_configs_dist.py

Shard s lives on rank s % N; every query goes through pilosa_b200.executor.Executor, i.e. one
C-ABI call per rank plus the NCCL merge of the partials (sum all-reduce for Count / Sum, gathers for Min / Max /
TopN). Per query rank 0 prints one JSON line: wall ms of the whole distributed call (median of the per-call
maximum over ranks), the local kernel ms of rank 0, and the RESULT, so that runs at different N (N = 1 included:
plain `python bench_configs_dist.py`) can be compared for equality — the full-size parity check across GPU counts.
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

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from pilosa_b200 import abi, synth  # noqa: E402
from pilosa_b200.abi import Difference, Intersect, Row, Union, Xor  # noqa: E402
from pilosa_b200.engine import Engine  # noqa: E402
from pilosa_b200.executor import Executor  # noqa: E402

SEED = 20190416


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="c2,c3,c4,c5")
    ap.add_argument("--reps", type=int, default=7)
    ap.add_argument("--c2-shards", type=int, default=1024)
    ap.add_argument("--c3-shards", type=int, default=1024)
    ap.add_argument("--c3-rows", type=int, default=10000)
    ap.add_argument("--c4-shards", type=int, default=2048)
    ap.add_argument("--c5-shards", type=int, default=8192)
    args = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = Engine(local)
    g = Executor(eng, rank=rank, world=world)
    cores = max(1, (os.cpu_count() or 8) // world)
    cfgs = args.configs.split(",")

    def load(field, gen, shards):
        mine = [s for s in shards if s % world == rank]
        with ThreadPoolExecutor(max_workers=cores) as tp:
            for s, fb in zip(mine, tp.map(gen, mine)):
                g.load_fragment(field, s, fb)

    def run(config, name, fn, shards, reps=None):
        reps = reps or args.reps
        fn(); fn()
        walls, kern = [], []
        for _ in range(reps):
            if world > 1:
                dist.barrier()
            t0 = time.perf_counter()
            res = fn()
            w = (time.perf_counter() - t0) * 1e3
            if world > 1:
                t = torch.tensor([w], dtype=torch.float64, device="cuda")
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                w = float(t.item())
            walls.append(w)
            kern.append(eng.last_timing().kernel_ms)
        if rank == 0:
            print(json.dumps({"config": config, "query": name, "n_gpus": world, "shards": len(shards), "call_ms_max_over_ranks": round(statistics.median(walls), 4),
                              "kernel_ms_rank0": round(statistics.median(kern), 4), "result": res}), flush=True)

    if "c2" in cfgs:
        shards = tuple(range(args.c2_shards))
        g.create_field("m")
        load("m", lambda s: synth.mixed(s, 0, 4, SEED), shards)
        leaves = [Row("m", r) for r in range(4)]
        for qname, call in (("Count(Union(r0..r3))", Union(*leaves)), ("Count(Difference(r0,r1,r2,r3))", Difference(*leaves)),
                            ("Count(Xor(r0..r3))", Xor(*leaves))):
            run("C2", qname, lambda: g.count(call, shards), shards)
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
            run("C4", f"Count(Row(v > {x}))", lambda x=x: g.count(g.row_range("v", abi.BSI_GT, x), shards), shards)
    if "c5" in cfgs:
        shards = tuple(range(args.c5_shards))
        g.create_field("sp")
        load("sp", lambda s: synth.bernoulli(s, 1, 2, 1e-4, SEED), shards)
        call = Intersect(Row("sp", 1), Row("sp", 2))
        for n in (64, 512, args.c5_shards):
            sh = shards[:n]
            run("C5", f"Count(Intersect(Row(sp=1),Row(sp=2))) p=1e-4, {n} shards", lambda: g.count(call, sh), sh)
            run("C5", f"batch of 256 x same query, {n} shards (whole batch)", lambda: g.count_batch([call] * 256, sh)[:2], sh)
    eng.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
