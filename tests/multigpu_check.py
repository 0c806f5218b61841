#!/usr/bin/env python
"""Multi-GPU parity check, run under torchrun on N >= 2 GPUs of one box:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29533 tests/multigpu_check.py

Every rank owns the shards s with s % N == rank in its own engine context; Count / Sum /
Min / Max / TopN / range counts are merged over NCCL by pilosa_b200.executor.Executor and
compared, on every rank, with the CPU oracle holding ALL shards. Not collected by pytest
(the driver's GPU tier has one GPU); the log of a run is kept under profiles/.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import oracle  # noqa: E402  (checker)
from pilosa_b200 import abi, synth  # noqa: E402
from pilosa_b200.abi import Difference, Intersect, Row, Union, Xor  # noqa: E402
from pilosa_b200.engine import Engine  # noqa: E402
from pilosa_b200.executor import Executor  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = Engine(local)
    g = Executor(eng, rank=rank, world=world)
    c = Executor(oracle.Oracle(workers=4))  # all shards, single "rank"
    shards = list(range(11)) + [20]  # uneven split across ranks; shard 20 exists only in some fields
    for ex in (g, c):
        ex.create_field("f")
        ex.create_field("m")
        ex.create_field("z")
        ex.create_field("v", type="int", min=-(1 << 20), max=(1 << 20) - 1)
    for s in shards:
        frs = {"f": synth.bernoulli(s, 0, 4, 0.10), "m": synth.mixed(s, 0, 4), "z": synth.zipf(s, n_rows=400, k=4),
               "v": synth.bsi(s, bit_depth=21, vmin=-(1 << 20), vmax=(1 << 20) - 1, density=0.3)}
        for name, fb in frs.items():
            if s == 20 and name in ("z",):
                continue
            g.load_fragment(name, s, fb)  # no-op on ranks that do not own s
            c.load_fragment(name, s, fb)
    src = Intersect(Row("f", 1), Row("f", 2))
    checks = []
    for call in (src, Union(Row("m", 0), Row("m", 1), Row("m", 2), Row("m", 3)), Difference(Row("m", 0), Row("f", 1), Row("m", 2)),
                 Xor(Row("m", 0), Row("m", 1), Row("f", 3))):
        checks.append((repr(call), g.count(call, shards), c.count(call, shards)))
    checks.append(("count_batch", g.count_batch([src, Row("m", 1)], shards), c.count_batch([src, Row("m", 1)], shards)))
    for filt in (None, Row("f", 1)):
        checks.append((f"sum/{filt}", g.sum("v", shards, filt), c.sum("v", shards, filt)))
        checks.append((f"min/{filt}", g.min("v", shards, filt), c.min("v", shards, filt)))
        checks.append((f"max/{filt}", g.max("v", shards, filt), c.max("v", shards, filt)))
    for x in (0, 12345, -777):
        checks.append((f"count(v>{x})", g.count(g.row_range("v", abi.BSI_GT, x), shards), c.count(c.row_range("v", abi.BSI_GT, x), shards)))
    for kw in (dict(n=10), dict(n=10, src=src), dict(n=5, src=Row("m", 1), threshold=3), dict(n=10, src=src, ids=[1, 2, 3, 399])):
        checks.append((f"topn/{kw}", g.topn("z", shards, **kw), c.topn("z", shards, **kw)))
    from pilosa_b200.executor import RowsCall as R
    for kw in (dict(), dict(limit=3), dict(previous=1, limit=2), dict(column=(3 << 20) + 17)):
        checks.append((f"rows/{kw}", g.rows("z", shards, **kw), c.rows("z", shards, **kw)))
    for ch, kw in (([R("f"), R("m")], {}), ([R("f"), R("m")], dict(filter_call=src, limit=5)), ([R("z", limit=6), R("f")], {}),
                   ([R("f", previous=1), R("m", previous=2)], {}), ([R("f"), R("m"), R("f")], dict(limit=20))):
        checks.append((f"groupby/{ch}/{kw}", g.group_by(ch, shards, **kw), c.group_by(ch, shards, **kw)))
    # bitmap-returning calls stay sharded: every rank checks its own segments against the oracle's
    from pilosa_b200.abi import Shift
    for call in (src, Union(Row("m", 0), Row("f", 1)), Shift(Row("f", 1), n=1), g.row_range("v", abi.BSI_LT, -5)):
        mine = g.row_segments(call, shards)
        exp = [(s, b) for s, b in c.row_segments(call, shards) if s % world == rank]
        checks.append((f"segments/{call!r}", mine, exp))
    # the protobuf Columns field of the rank-local part of a result (serialized Row messages concatenate: the
    # coordinator appends the ranks' bytes), and a Union wider than one device program, counted across ranks
    for call in (src, Shift(Row("f", 1), n=1), g.row_range("v", abi.BSI_GT, 1 << 19)):
        checks.append((f"proto/{call!r}", g.row_proto_columns(call, shards),
                       c.be.eval_proto_columns(c.program(call), [s for s in sorted(shards) if s % world == rank])))
    wide = Union(*[Row("z", r) for r in range(100)])
    checks.append(("wide union", g.count(wide, shards), c.count(wide, shards)))
    checks.append(("wide in a batch", g.count_batch([wide, src, Difference(Row("m", 0), *[Row("z", r) for r in range(40)])], shards),
                   c.count_batch([wide, src, Difference(Row("m", 0), *[Row("z", r) for r in range(40)])], shards)))
    # counts summed over the ranks by the batch's own kernel (peer mailboxes over NVLink, no NCCL): prepared batches
    # launched back to back, several epochs deep, and the split _begin / _end form
    import ctypes as C

    handles = [None] * world
    dist.all_gather_object(handles, eng.peer_export())
    eng.peer_connect(rank, world, b"".join(handles))
    local = [s for s in shards if s % world == rank]
    qsets = [[src, Row("m", 1), Union(Row("m", 0), Row("f", 3))], [Row("f", 0), Xor(Row("m", 2), Row("m", 3))], [Intersect(Row("m", 0), Row("f", 1), Row("f", 2))] * 5]
    sh = abi.u64_array(local)
    prepared = []
    for qs in qsets:
        flat, off = eng.pack_programs([g.program(q) for q in qs])
        prepared.append((eng.batch_prepare(flat, off, len(qs), sh, len(local)), qs, flat, off))
    for rep in range(3):                      # 9 launches in a row: the mailbox slots are reused twice
        for b, _, _, _ in prepared:
            eng.batch_launch_allreduce(b)
    for b, qs, _, _ in prepared:
        out = (C.c_uint64 * len(qs))()
        eng.batch_fetch(b, out)
        checks.append((f"peer reduce / prepared x{len(qs)}", list(out), [c.count(q, shards) for q in qs]))
    eng.set_option(eng.OPT_PEER_ALLREDUCE, 1)
    outs = []
    for k in range(5):                        # two batches in flight, like bench.py's e2e leg
        _, qs, flat, off = prepared[k % 3]
        eng.count_batch_begin(flat, off, len(qs), sh, len(local))
        if k:
            pq = prepared[(k - 1) % 3][1]
            out = (C.c_uint64 * len(pq))()
            eng.count_batch_end(out)
            outs.append((list(out), [c.count(q, shards) for q in pq]))
    pq = prepared[4 % 3][1]
    out = (C.c_uint64 * len(pq))()
    eng.count_batch_end(out)
    outs.append((list(out), [c.count(q, shards) for q in pq]))
    eng.set_option(eng.OPT_PEER_ALLREDUCE, 0)
    checks.append(("peer reduce / begin-end", [a for a, _ in outs], [b for _, b in outs]))
    for b, _, _, _ in prepared:
        eng.batch_free(b)
    bad = [(n, a, b) for n, a, b in checks if a != b]
    ok = torch.tensor([0 if bad else 1], device="cuda")
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    if bad:
        print(f"rank {rank}: MISMATCH {bad}", flush=True)
    if rank == 0:
        print(f"MULTIGPU {'OK' if int(ok.item()) else 'FAILED'}: {len(checks)} checks x {world} ranks, local shards per rank "
              f"{[len([s for s in shards if s % world == r]) for r in range(world)]}", flush=True)
    eng.close()
    dist.destroy_process_group()
    sys.exit(0 if int(ok.item()) else 1)


if __name__ == "__main__":
    main()
