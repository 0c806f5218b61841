#!/usr/bin/env python
"""bench.py — Count(Intersect(Row(f=a), Row(f=b))) throughput of the B200 engine.

Workload (BASELINE.json configs[0] scaled to where HBM bandwidth is measurable, SURVEY.md
§8d "C1 batched"): index `i`, set field `f`, 16 shards x 2^20 columns PER GPU, R rows of iid
Bernoulli(0.10) columns (every container is a bitmap container), one step = one batch of B
distinct random row pairs (a, b), each a Count(Intersect(Row(f=a), Row(f=b))) over all
shards. Per step every GPU reads B x 16 shards x 16 containers x 2 x 8 KiB = 4 GiB of
operand payload, far beyond the 126 MB L2, so no L2 flush is needed between steps.

  value   device-resident throughput: prepared batches (programs already in HBM) launched
          back to back, timed with CUDA events on the launching stream, max over ranks.
  e2e     the same metric through the C-ABI call with HOST buffers: per step the query
          programs are copied host->device, the kernel runs and the counts come back
          device->host (pb200_count_batch_begin/_end), wall clock, max over ranks.
  N > 1   shards are partitioned over ranks (shard s -> rank s % N, 16 shards per GPU, weak
          scaling); the only exchange is the all-reduce(sum) of the B per-query counts over
          NCCL, inside the timed step. `value` counts 16-shard query units: N x B per step.

--impl reference times the CPU restatement of the reference's Go path (oracle/, mode A:
materialise the intersection, then count — executor.go:1804-1810) on the host cores.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import random
import statistics
import subprocess
import sys
import threading
import time
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SHARDS_PER_GPU = 16
P_DENSITY = 0.10
SEED = 20190416
METRIC = "count_intersect_queries_per_sec"
UNIT = "queries/s"
CPU_SAMPLE_ROWS = 1024  # rows the CPU leg of the native arm holds: 16 shards x 1024 rows x 128 KiB = 2 GiB, far beyond any LLC


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index: int):
        self.idx = gpu_index
        self.proc = None
        self.lines = []
        self.nvml = None
        self._stop = False
        self._samples = []

    def _nvml_loop(self):
        import pynvml as nv

        h = nv.nvmlDeviceGetHandleByIndex(self.idx)
        bits = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown, "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown, "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
        mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
        while not self._stop:
            sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
            r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
            try:
                pw = nv.nvmlDeviceGetPowerUsage(h) / 1000.0
            except Exception:
                pw = None
            self._samples.append((sm, mx, [k for k, b in bits.items() if r & b], pw, time.time()))
            time.sleep(0.002)

    def start(self):
        try:  # NVML polling every ~2 ms: the timed region is only tens of milliseconds long
            import pynvml as nv

            nv.nvmlInit()
            self.nvml = threading.Thread(target=self._nvml_loop, daemon=True)
            self.nvml.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(self.idx)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def mark(self):
        """Start of the timed region: earlier samples (the poller is started BEFORE the barrier so that rank 0 does not
        enter the region late) are dropped when the result is formed."""
        self.t_mark = time.time()

    def stop(self):
        if self.nvml is not None:
            self._stop = True
            self.nvml.join(timeout=2)
            t_mark = getattr(self, "t_mark", 0.0)
            kept = [s for s in self._samples if s[4] >= t_mark]
            self._samples = kept or self._samples[-1:]
            sm = [s[0] for s in self._samples]
            reasons = sorted({r for s in self._samples for r in s[2]})
            pw = [s[3] for s in self._samples if s[3] is not None]
            return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": self._samples[0][1] if self._samples else None,
                    "sm_mhz_min": min(sm) if sm else None, "power_w_max": max(pw) if pw else None,
                    "reasons": reasons, "samples": len(sm), "source": "nvml, 2 ms period, timed region only"}
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


class DevArray:
    """Zero-copy view of a device buffer for torch (CUDA array interface)."""

    def __init__(self, ptr: int, n: int):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<i8", "data": (ptr, False), "version": 2}


def make_batches(n_batches: int, batch: int, rows: int):
    rnd = random.Random(SEED)
    out = []
    for _ in range(n_batches):
        seen, pairs = set(), []
        while len(pairs) < batch:
            a, b = rnd.randrange(rows), rnd.randrange(rows)
            if a != b and (a, b) not in seen:
                seen.add((a, b))
                pairs.append((a, b))
        out.append(pairs)
    return out


def pack_pairs(pairs, fam: int):
    """(flat pb200_node[3B], prog_off[B+1]) for Count(Intersect(Row(a), Row(b))) queries."""
    from pilosa_b200 import abi

    n = len(pairs)
    flat = (abi.Node * (3 * n))()
    off = (C.c_uint32 * (n + 1))()
    for i, (a, b) in enumerate(pairs):
        flat[3 * i] = abi.Node(op=abi.OP_ROW, family=fam, row=a)
        flat[3 * i + 1] = abi.Node(op=abi.OP_ROW, family=fam, row=b)
        flat[3 * i + 2] = abi.Node(op=abi.OP_INTERSECT, nargs=2)
        off[i] = 3 * i
    off[n] = 3 * n
    return flat, off


def mem_available_bytes() -> int:
    try:
        for ln in open("/proc/meminfo"):
            if ln.startswith("MemAvailable:"):
                return int(ln.split()[1]) * 1024
    except OSError:
        pass
    return 0


class CpuArm:
    """The reference's CPU path for this workload, restated (oracle/): an index over rows [0, rows_held) of the same
    per-(shard, row, container) streams the GPU arm loads, a PERSISTENT worker pool of `cores` threads (executor.go:83-107),
    and one step = the whole batch of Count(Intersect(Row, Row)) calls in flight together, i.e. batch x shards per-shard
    jobs on that pool (what concurrent requests give the reference; executor.go:2576-2611).
      mode A  executeCount as the reference runs it: materialise the intersection, then count (executor.go:1804-1810)
      mode B  best-effort CPU: fused intersectionCount, nothing materialised (roaring.go:568-590)"""

    def __init__(self, shards, rows_held: int, batch: int, cores: int):
        import oracle
        from pilosa_b200 import synth

        self.shards, self.rows_held, self.batch, self.cores = list(shards), rows_held, batch, cores
        self.orc = oracle.Oracle(workers=cores)
        self.fam = self.orc.family_create()
        t0 = time.perf_counter()

        def gen(s):
            fb = synth.bernoulli(s, 0, rows_held, P_DENSITY, SEED)
            self.orc.frag_load(self.fam, s, fb)   # ctypes releases the GIL: shards are generated and parsed in parallel

        with ThreadPoolExecutor(max_workers=max(1, min(cores, 8, len(self.shards)))) as tp:
            list(tp.map(gen, self.shards))
        self.build_s = time.perf_counter() - t0
        self.sh = (C.c_uint64 * len(self.shards))(*self.shards)
        self.index_bytes = len(self.shards) * rows_held * 16 * 8192

    def step(self, pairs, fused: bool):
        flat, off = pack_pairs(pairs, self.fam)
        return self.orc.count_batch_raw(flat, off, len(pairs), self.sh, len(self.shards), fused)

    def time_mode(self, batches, fused: bool, steps: int, min_seconds: float):
        """(queries/s, steps run, seconds). Batches are cycled; packing the programs is outside the clock."""
        packed = [pack_pairs(p, self.fam) for p in batches]
        n, t0 = 0, time.perf_counter()
        while n < steps or (time.perf_counter() - t0) < min_seconds:
            fl, off = packed[n % len(packed)]
            self.orc.count_batch_raw(fl, off, self.batch, self.sh, len(self.shards), fused)
            n += 1
        dt = time.perf_counter() - t0
        return n * self.batch / dt, n, dt

    def describe(self, qa, qb, n_steps, secs) -> dict:
        per_q = len(self.shards) * 16 * 2 * (8192 + 16)
        return {"value": qa, "unit": UNIT, "cores": self.cores, "kind": "port",
                "sample": f"{self.batch} queries per step ({self.batch * len(self.shards)} per-shard jobs on a persistent pool of {self.cores} "
                          f"threads) over rows [0,{self.rows_held}) x {len(self.shards)} shards = {self.index_bytes / 2**30:.1f} GiB held in RAM, "
                          f"{n_steps} steps in {secs:.1f} s; value = mode A (materialise then count, executor.go:1804)",
                "mode_a_materialise_then_count": qa, "mode_b_fused_intersection_count": qb,
                "mode_a_algorithmic_gb_per_s": qa * per_q / 1e9, "mode_b_algorithmic_gb_per_s": qb * per_q / 1e9,
                "index_build_s": self.build_s, "seconds": secs}


def cpu_rows_held(rows: int, shards: int, want_all: bool) -> int:
    """All rows of the GPU arm's index when the host has the memory for it (index + generator blobs), else a prefix."""
    if want_all:
        need = shards * rows * 16 * 8192
        if mem_available_bytes() > 2.2 * need + (8 << 30):
            return rows
    return min(rows, CPU_SAMPLE_ROWS)


def run_reference(args):
    """--impl reference: the CPU arm alone, on this arm's config. Each step is the full batch of the GPU arm (same
    generator, same seed, same pairs) when the whole index fits in host RAM; otherwise the batch is drawn over the
    prefix of rows that is held, and `sample` says so."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    shards = list(range(SHARDS_PER_GPU))
    held = cpu_rows_held(args.rows, len(shards), True)
    arm = CpuArm(shards, held, args.batch, cores)
    batches = make_batches(args.warmup + args.steps, args.batch, held)
    for i in range(args.warmup):
        arm.step(batches[i], False)
    qa, n, secs = arm.time_mode(batches[args.warmup:], False, args.steps, 0.0)
    qb, _, _ = arm.time_mode(batches[args.warmup:], True, max(1, args.steps // 2), 0.0)
    cpu = arm.describe(qa, qb, n, secs)
    if args.cpu_scan:  # throughput against pool size: the pool, not thread start-up, is what is being timed
        scan = {}
        for w in sorted({1, 2, 4, max(1, cores // 2), cores}):
            arm.orc.set_workers(w)
            a, _, _ = arm.time_mode(batches[args.warmup:], False, 2, 0.0)
            b, _, _ = arm.time_mode(batches[args.warmup:], True, 2, 0.0)
            scan[str(w)] = {"mode_a": a, "mode_b": b}
        cpu["workers_scan"] = scan
    line = {
        "impl": "reference", "metric": METRIC, "value": qa, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * args.batch / qa, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": workload_config(args, 1),
        "cpu_baseline": cpu,
        "e2e": {"value": qa, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "reference_arm": "CPU restatement of the reference's Go path (oracle/; the Go toolchain is not in this image), all host cores",
    }
    print(json.dumps(line), flush=True)


def workload_config(args, world):
    return {
        "workload": f"C1 batched: Count(Intersect(Row(f=a),Row(f=b))), {SHARDS_PER_GPU} shards x 2^20 cols per GPU, 10% random density "
                    f"(all bitmap containers), {args.rows} rows, batch of {args.batch} distinct random row pairs per step",
        "shards_per_gpu": SHARDS_PER_GPU, "total_shards": SHARDS_PER_GPU * world, "rows": args.rows, "batch": args.batch,
        "density": P_DENSITY, "seed": SEED, "partition": f"shard s -> rank s % {world}",
        "l2": "inputs larger than L2 (4 GiB of operands per step per GPU); no flush needed",
        "value_unit_note": "one unit = one Count(Intersect) over 16 shards (16.8M columns per operand row); a query over the whole "
                           "N x 16-shard index is N units",
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--rows", type=int, default=int(os.environ.get("PB200_BENCH_ROWS", "4096")))
    ap.add_argument("--batch", type=int, default=1024)
    ap.add_argument("--cpu-seconds", type=float, default=10.0)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--cpu-scan", action="store_true", help="reference arm: also time the pool at 1/2/4/P workers")
    ap.add_argument("--sustained-seconds", type=float, default=2.0)
    ap.add_argument("--no-strong", action="store_true", help="skip the strong-scaling leg (fixed 128-shard index)")
    ap.add_argument("--trace-steps", action="store_true", help="stderr: device time of every timed step of the resident leg (diagnostic)")
    args = ap.parse_args()
    if args.warmup < 3:
        raise SystemExit("bench.py: --warmup must be >= 3 (timing rule); got %d" % args.warmup)

    if args.impl == "reference":
        run_reference(args)
        return

    import torch
    import torch.distributed as dist

    from pilosa_b200 import synth
    from pilosa_b200.engine import Engine

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: pilosa_b200 has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    stream = torch.cuda.Stream()
    eng = Engine(local_rank)
    eng.set_stream(stream.cuda_stream)
    fam = eng.family_create()
    shards = [rank + world * i for i in range(SHARDS_PER_GPU)]
    if world > 1:  # the ranks' mailboxes are exchanged once; from then on the sum over the ranks is the tail of each batch's kernel
        handles = [None] * world
        dist.all_gather_object(handles, eng.peer_export())
        eng.peer_connect(rank, world, b"".join(handles))

    # ---- build + load the index (not timed) ----
    t_build = time.perf_counter()
    workers = max(1, min(8, (os.cpu_count() or 8) // max(1, world)))
    with ThreadPoolExecutor(max_workers=workers) as tp:
        futs = [tp.submit(synth.bernoulli, s, 0, args.rows, P_DENSITY, SEED) for s in shards]
        for s, f in zip(shards, futs):
            fb = f.result()
            eng.frag_load(fam, s, fb)
            del fb
    t_build = time.perf_counter() - t_build
    shards_arr = (C.c_uint64 * len(shards))(*shards)

    n_batches = args.warmup + args.steps
    batches = [pack_pairs(p, fam) for p in make_batches(n_batches, args.batch, args.rows)]
    B = args.batch

    def barrier():
        """All ranks idle, then released together: the processes of one box share a clock, so after the collective
        barrier rank 0 names a start time 2 ms ahead and every rank spins up to it — the ranks leave an NCCL barrier
        hundreds of microseconds apart, which a 20-step timed region would otherwise carry as its tail."""
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            t = torch.tensor([time.time() + 0.002], dtype=torch.float64, device="cuda")
            dist.broadcast(t, src=0)
            deadline = float(t.item())
            torch.cuda.synchronize()
            while time.time() < deadline:
                pass
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    sampler = ClockSampler(local_rank)
    with torch.cuda.stream(stream):
        # ---- leg 1: device-resident (prepared batches) ----
        prepared = [eng.batch_prepare(fl, off, B, shards_arr, len(shards)) for fl, off in batches]

        # N > 1: the sum of a batch's counts over the ranks is done by the batch's own kernel — its last CTA adds this GPU's
        # counts into every rank's mailbox over NVLink peer memory (pb200_batch_launch_allreduce); no collective is launched
        # between two batches. A batch's total is picked up two launches later in stream order (join_side flushes the rest).
        def step_resident(i):
            if world > 1:
                eng.batch_launch_allreduce(prepared[i])
            else:
                eng.batch_launch(prepared[i])

        def join_side():
            if world > 1:
                eng.peer_flush()

        for i in range(args.warmup):
            step_resident(i)
        join_side()
        if rank == 0:  # one poller per box: concurrent NVML queries from every rank serialise in the driver
            sampler.start()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        sampler.mark()
        ev0.record(stream)
        marks = []
        for i in range(args.warmup, n_batches):
            step_resident(i)
            if args.trace_steps:
                marks.append(torch.cuda.Event(enable_timing=True))
                marks[-1].record(stream)
        join_side()                      # the last all-reduce is inside the timed region
        ev1.record(stream)
        barrier()
        ms_total = max_over_ranks(ev0.elapsed_time(ev1))
        if args.trace_steps:
            ts = [ev0.elapsed_time(m) for m in marks] + [ev0.elapsed_time(ev1)]
            print(f"rank {rank} step ends (ms): " + " ".join(f"{t:.3f}" for t in ts), file=sys.stderr, flush=True)
        out = (C.c_uint64 * B)()
        eng.batch_fetch(prepared[-1], out)
        last_resident = list(out)

        # ---- leg 1b: the same resident loop held for >= --sustained-seconds (clocks and power settle; the 20-step
        # figure above is a burst) ----
        sustained = None
        if args.sustained_seconds > 0:
            n_sus = max(args.steps, int(args.sustained_seconds / max(1e-6, ms_total / args.steps * 1e-3)) + 1)
            sus_sampler = ClockSampler(local_rank)
            if rank == 0:
                sus_sampler.start()
            s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier()
            sus_sampler.mark()
            s0.record(stream)
            for k in range(n_sus):
                step_resident(k % len(prepared))
            join_side()
            s1.record(stream)
            barrier()
            sus_ms = max_over_ranks(s0.elapsed_time(s1))
            sus_clocks = sus_sampler.stop() if rank == 0 else None
            sustained = {"steps": n_sus, "seconds": sus_ms * 1e-3, "ms_per_step": sus_ms / n_sus,
                         "value": B * world / (sus_ms / n_sus * 1e-3), "unit": UNIT, "clocks": sus_clocks}
        for b in prepared:
            eng.batch_free(b)

        # ---- leg 2: end to end through the C ABI with host buffers ----
        out_e2e = (C.c_uint64 * B)()

        def begin(i):
            fl, off = batches[i]
            eng.count_batch_begin(fl, off, B, shards_arr, len(shards))   # host programs -> pinned -> H2D, kernels (async)

        def run_e2e(idx):
            """Every step: programs host->device, kernel, counts device->host. On one GPU the host
            prepares step i+1 while step i runs (two batches in flight); with N GPUs the NCCL
            all-reduce of the device-resident counts is queued right behind each batch's kernel."""
            if world == 1:
                begin(idx[0])
                for n in range(len(idx)):
                    if n + 1 < len(idx):
                        begin(idx[n + 1])
                    eng.count_batch_end(out_e2e)                           # D2H on the copy stream + wait
            else:  # PB200_OPT_PEER_ALLREDUCE: _begin's kernel pushes to the peers, _end waits for their counts on the copy stream
                begin(idx[0])
                for n in range(len(idx)):
                    if n + 1 < len(idx):
                        begin(idx[n + 1])
                    eng.count_batch_end(out_e2e)

        if world > 1:
            eng.set_option(eng.OPT_PEER_ALLREDUCE, 1)
        run_e2e(list(range(args.warmup)))
        barrier()
        t0 = time.perf_counter()
        run_e2e(list(range(args.warmup, n_batches)))
        torch.cuda.synchronize()
        e2e_s = max_over_ranks(time.perf_counter() - t0)
        barrier()
        clocks = sampler.stop() if rank == 0 else None
        tm = eng.last_timing()
        assert list(out_e2e) == last_resident, "resident and e2e legs disagree"

        # ---- leg 3: strong scaling on a FIXED index — 128 shards x 512 rows (8 GiB), whatever N is; rank r holds the
        # shards s < 128 with s % N == r, so the per-GPU work shrinks as N grows. One step = the same 1024-pair batch over
        # the whole 128-shard index (+ the all-reduce of the counts for N > 1). ----
        strong = None
        if not args.no_strong:
            S_SHARDS, S_ROWS = 128, 512
            fam2 = eng.family_create()
            mine = [s for s in range(S_SHARDS) if s % world == rank]
            with ThreadPoolExecutor(max_workers=workers) as tp:
                futs = [tp.submit(synth.bernoulli, s, 0, S_ROWS, P_DENSITY, SEED) for s in mine]
                for s, f in zip(mine, futs):
                    eng.frag_load(fam2, s, f.result())
            mine_arr = (C.c_uint64 * len(mine))(*mine)
            sbatches = [pack_pairs(pp, fam2) for pp in make_batches(n_batches, B, S_ROWS)]
            sprep = [eng.batch_prepare(fl, off, B, mine_arr, len(mine)) for fl, off in sbatches]

            def step_strong(i):
                if world > 1:
                    eng.batch_launch_allreduce(sprep[i])
                else:
                    eng.batch_launch(sprep[i])

            for i in range(args.warmup):
                step_strong(i)
            join_side()
            barrier()
            q0, q1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            q0.record(stream)
            for i in range(args.warmup, n_batches):
                step_strong(i)
            join_side()
            q1.record(stream)
            barrier()
            s_ms = max_over_ranks(q0.elapsed_time(q1)) / args.steps
            chk = (C.c_uint64 * B)()
            eng.batch_fetch(sprep[-1], chk)
            strong = {"index": f"{S_SHARDS} shards x {S_ROWS} rows x 2^20 columns, 10% density ({S_SHARDS * S_ROWS * 16 * 8192 / 2**30:.0f} GiB), fixed for every N",
                      "shards_per_gpu": len(mine), "ms_per_step": s_ms, "value": B / (s_ms * 1e-3), "unit": "queries/s over the whole fixed index",
                      "checksum_last_batch": int(sum(chk) % (1 << 61)),
                      "note": "strong scaling: divide this value at N by the value at N = 1 (the checksum must be equal at every N)"}
            for b in sprep:
                eng.batch_free(b)

    ms_per_step = ms_total / args.steps
    units_per_step = B * world
    value = units_per_step / (ms_per_step * 1e-3)
    e2e_value = units_per_step * args.steps / e2e_s
    algo_bytes = B * len(shards) * 16 * 2 * (8192 + 16)   # per launch per GPU (payload + 16 B/container)
    peak, peak_src = peaks()
    achieved = algo_bytes / (ms_per_step * 1e-3) / 1e9
    traffic, traffic_src = None, None
    try:  # dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu --set full capture
        prof = json.load(open(os.path.join(ROOT, "profiles", "r02_eval_fast_summary.json")))
        if B == 1024 and len(shards) == SHARDS_PER_GPU and args.rows == 4096:  # the capture was taken on exactly this configuration
            traffic, traffic_src = prof["traffic_bytes_per_launch"], "ncu capture of this configuration, not of this run: profiles/r02_eval_fast_summary.json (" + prof["command"] + ")"
    except Exception:
        pass
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64",
        "data": "synthetic", "config": workload_config(args, world),
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(tm.h2d_bytes), "d2h_bytes_per_step": int(tm.d2h_bytes) + 8,
                "ms_per_step": 1e3 * e2e_s / args.steps},
        "gpu_launches": args.steps * 2 * (1 + (1 if world > 1 else 0)),  # timed steps of both legs: one eval_fast_kernel each (+ one pick-up kernel for N > 1)
        "clocks": clocks,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                     "kernel": "eval_fast_kernel (K1: 128-bit AND + popcount over bitmap container pairs)",
                     "algo_bytes_per_launch": algo_bytes, "peak_source": peak_src,
                     "note": "achieved = algorithmic bytes per launch / mean launch duration from CUDA events over the timed region "
                             "(includes the per-step 8 KiB count memset" + (" and the peer reduction of the counts in the kernel's tail" if world > 1 else "") + ")"},
        "sustained": sustained,
        "strong_scaling": strong,
        "columns_per_s": value * SHARDS_PER_GPU * (1 << 20) * 2,
        "global_queries_per_s": value / world,
        "index_build_s": t_build,
        "engine_algo_bytes_last_call": int(tm.algo_bytes),
    }

    # ---- CPU baseline (rank 0, N = 1 only) + parity check against it ----
    if rank == 0 and world == 1 and not args.no_cpu:
        cores = os.cpu_count() or 1
        held = cpu_rows_held(args.rows, len(shards), False)
        arm = CpuArm(shards, held, args.batch, cores)
        cb = make_batches(4, args.batch, held)
        cpu_counts = list(arm.step(cb[0], False))           # warm-up step; its counts are the parity reference
        qa, n_a, secs_a = arm.time_mode(cb, False, 1, args.cpu_seconds * 0.6)
        qb, _, _ = arm.time_mode(cb, True, 1, args.cpu_seconds * 0.4)
        fl, off = pack_pairs(cb[0], fam)
        got = (C.c_uint64 * args.batch)()
        eng.count_batch_raw(fl, off, args.batch, shards_arr, len(shards), got)
        ok = list(got) == cpu_counts
        line["cpu_baseline"] = arm.describe(qa, qb, n_a, secs_a)
        line["parity_check"] = "ok: %d queries (one full batch) bit-exact vs oracle" % args.batch if ok else "MISMATCH vs oracle"
        del arm
        if not ok:
            print(json.dumps(line), flush=True)
            raise SystemExit("parity check against the oracle failed")
    if rank == 0:
        print(json.dumps(line), flush=True)
    eng.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
