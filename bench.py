#!/usr/bin/env python
"""bench.py -- word-count tokens/s (map + shuffle + reduce) on B200, BASELINE.json's metric.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

Headline workload: BASELINE.json configs[1], the synthetic Europarl-shaped corpus (197 splits x
10,000 lines, ~49 M Zipf tokens, R = 10).  With N ranks every rank maps its own 197-split shard
(weak scaling) and the partitions are merged through one exchange.  A step = one whole job over the
resident splits: tokenise + combine, partition, (shuffle,) sort, segmented reduce, rows left in HBM.

Beside it, in the same JSON line:
  parity_check    the first (untimed) job of every workload is compared with the CPU oracle run on
                  the union of all ranks' splits: an order-independent 64-bit digest over every
                  (partition, key, count) row, the row count, the count sum, and the order of the rows
                  on each rank; a mismatch ends the run with a non-zero exit code.
  strong_config3  BASELINE.json configs[2]: the FIXED 1,970-split corpus (~491 M tokens) sharded
                  round-robin over the N ranks (one map job per split, server.lua:256-272), so that
                  t_1 / (N * t_N) over the N = 1, 2, 4, 8 lines is the strong-scaling efficiency
                  SURVEY.md 8(d) defines.
Prints ONE JSON line (rank 0).
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "wordcount_tokens_per_s"
UNIT = "tokens/s"
R = 10
SEED = 0x5EED0001
CONFIG3_SCALE = 10


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scale", type=int, default=1, help="corpus scale per rank (1 = config 2 shape)")
    ap.add_argument("--shuffle", default="auto", choices=["auto", "p2p", "nccl"],
                    help="N>1: peer-to-peer windows (auto: where the GPUs can map each other) or NCCL all-to-all")
    ap.add_argument("--mod-owners", action="store_true", help="N>1: owner = partition mod ranks instead of balanced key ranges")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-strong", action="store_true", help="skip the config-3 strong-scaling block")
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle comparison of the first job")
    return ap.parse_args()


def host_threads():
    """Threads the CPU arms use: every core this process may run on.  Passed to the oracle explicitly
    (its OpenMP regions carry num_threads), so that OMP_NUM_THREADS=1, which torch.distributed.run
    injects into every rank, does not turn the CPU arm into a single-core run."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock and throttle reasons polled through NVML during the timed region (the steps are
    far shorter than nvidia-smi's sampling period, so the recipe's query is done in-process)."""

    def __init__(self, index):
        self.index, self.sm, self.reasons, self.stop_flag, self.max = index, [], 0, False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _poll(self):
        nv = self.nv
        while not self.stop_flag:
            try:
                self.sm.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                self.reasons |= nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            except Exception:
                pass
            time.sleep(0.002)

    def start(self):
        if self.nv:
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()

    def stop(self):
        if not self.nv:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable"]}
        self.stop_flag = True
        self.t.join(1)
        nv, names = self.nv, []
        for name, attr in (("hw_slowdown", "nvmlClocksThrottleReasonHwSlowdown"),
                           ("hw_thermal_slowdown", "nvmlClocksThrottleReasonHwThermalSlowdown"),
                           ("sw_thermal_slowdown", "nvmlClocksThrottleReasonSwThermalSlowdown"),
                           ("sw_power_cap", "nvmlClocksThrottleReasonSwPowerCap")):
            if self.reasons & getattr(nv, attr, 0):
                names.append(name)
        sm = sorted(self.sm)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.max, "samples": len(sm),
                "reasons": names}


# ---------------------------------------------------------------------------- parity digest
def rows_digest64(part, count, key_off, key_bytes):
    """Order-independent 64-bit digest of (partition, key, count) rows: the sum, modulo 2^64, of a mixed
    hash of every row.  Returns (digest, rows, sum of counts).  Equal multisets of rows give equal
    digests wherever the rows live, so the per-rank digests of a sharded job add up to the oracle's."""
    import numpy as np
    n = len(part)
    if n == 0:
        return 0, 0, 0
    u64 = np.uint64
    off = np.asarray(key_off, dtype=np.int64)
    lens = np.diff(off)
    assert lens.min() >= 1
    kb = np.asarray(key_bytes, dtype=np.uint8)[int(off[0]):int(off[-1])].astype(u64)
    pos = np.arange(len(kb), dtype=np.int64) - np.repeat(off[:-1] - off[0], lens)
    pw = np.empty(int(lens.max()), dtype=u64)
    x = 1
    for j in range(len(pw)):
        pw[j] = x
        x = (x * 0x100000001B3) & 0xFFFFFFFFFFFFFFFF
    with np.errstate(over="ignore"):
        kh = np.add.reduceat((kb + u64(1)) * pw[pos], (off[:-1] - off[0]))
        h = kh ^ (lens.astype(u64) * u64(0x9E3779B97F4A7C15))
        h = h * u64(0xD6E8FEB86659FD93) + np.asarray(part).astype(u64) * u64(0xA24BAED4963EE407) \
            + np.asarray(count).astype(u64) * u64(0x9FB21C651E98DF25)
        h ^= h >> u64(31)
        h *= u64(0xBF58476D1CE4E5B9)
        h ^= h >> u64(29)
        d = int(h.sum(dtype=u64))
        csum = int(np.asarray(count).astype(u64).sum(dtype=u64))
    return d, n, csum


def rows_in_order(rows):
    """partition ascending, key strictly ascending inside a partition (unsigned bytes, shorter first:
    Python's bytes order is Lua's string order in the C locale)."""
    keys = rows.keys()
    part = [int(p) for p in rows.part]
    return all((part[i], keys[i]) < (part[i + 1], keys[i + 1]) for i in range(len(keys) - 1))


class Env:
    pass


def bind_to_gpu_numa(index):
    """N > 1: run this rank on the CPUs next to its GPU (NVML's CPU affinity of the device) before any
    pinned buffer is allocated, so that first-touch places the staging memory on the GPU's own NUMA
    node and eight ranks' host-to-device copies do not all cross one memory controller.  Returns a
    short description for the JSON line."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1}
        cpus &= set(os.sched_getaffinity(0))
        if cpus:
            os.sched_setaffinity(0, cpus)
            return f"{len(cpus)} cpus {min(cpus)}-{max(cpus)}"
    except Exception as e:          # no NVML, no permission: keep the inherited affinity
        return f"unchanged ({type(e).__name__})"
    return "unchanged"


def parity_check(env, ctx, union_splits, info, balanced):
    """Every rank digests the rows it owns; the sums over ranks must equal the oracle's for the union
    corpus (run on rank 0 with all host threads).  Collective.  Returns the dict for the JSON line."""
    import numpy as np
    import torch
    import torch.distributed as dist
    rows = ctx.rows()
    d, n, csum = rows_digest64(rows.part, rows.count, rows.key_off, rows.key_bytes)
    ordered = rows_in_order(rows)
    keys = rows.keys()
    ends = None
    if n:
        ends = ((int(rows.part[0]), keys[0]), (int(rows.part[-1]), keys[-1]))
    mine = np.array([d, n, csum, int(info.n_tokens), int(ordered)], dtype=np.uint64).view(np.int64)
    t = torch.from_numpy(mine.copy()).to(env.dev)
    all_ends = [ends]
    if env.world > 1:
        dist.all_reduce(t)          # int64 sums wrap modulo 2^64, which is what the digest wants
        all_ends = [None] * env.world
        dist.all_gather_object(all_ends, ends)
    got = [int(v) & 0xFFFFFFFFFFFFFFFF for v in t.cpu().numpy().view(np.uint64)]
    verdict = [1]
    res = None
    if env.rank == 0:
        from oracle import oracle as O
        t0 = time.perf_counter()
        want = O.run_staged(union_splits(), R, host_threads())
        wd, wn, wsum = rows_digest64(want.part, want.count, want.key_off, want.key_bytes)
        ok = (got[0] == wd and got[1] == wn and got[2] == wsum == want.tokens and got[3] == want.tokens
              and got[4] == env.world)
        if balanced and env.world > 1:
            # key-range owners: rank r's rows all precede rank r+1's in the job's output order
            seq = [e for e in all_ends if e is not None]
            ok = ok and all(seq[i][1] < seq[i + 1][0] for i in range(len(seq) - 1))
        res = {"ok": bool(ok), "against": "oracle staged path (job.lua + utils.merge_iterator restated in C) over the union of all ranks' splits",
               "digest": f"{got[0]:016x}", "oracle_digest": f"{wd:016x}", "rows": got[1], "oracle_rows": wn,
               "count_sum": got[2], "tokens": want.tokens, "ranks_in_order": got[4], "oracle_s": round(time.perf_counter() - t0, 2)}
        verdict = [1 if ok else 0]
    if env.world > 1:
        dist.broadcast_object_list(verdict, src=0)
    if not verdict[0]:
        if env.rank == 0:
            print(json.dumps({"parity_check": res}), file=sys.stderr)
        raise SystemExit(f"bench.py: rank {env.rank}: rows differ from the oracle's -- not a valid run")
    return res


# ---------------------------------------------------------------------------- one workload on the GPUs
def run_workload(env, args, splits, tokens_local, union_splits, steps, warmup, want_e2e, pin):
    """Stages `splits` on this rank, checks the first job, then times `steps` resident jobs.
    Returns a dict of measurements (identical keys on every rank; aggregates are over all ranks)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from lua_mapreduce_1_b200 import _native, parallel
    world, rank, dev = env.world, env.rank, env.dev
    bytes_local = sum(len(s) for s in splits)
    max_keys = 1 << 20
    ctx = _native.Context(R, bytes_local + 64 * len(splits) + (1 << 16), max_keys=max_keys * (2 if world > 1 else 1),
                          table_slots=1 << 21, device=env.local_rank, stream=env.stream.cuda_stream, world_size=world, rank=rank)
    bufs = [torch.from_numpy(s).pin_memory() for s in splits] if pin else [torch.from_numpy(s) for s in splits]

    def stage_all():
        for i, t in enumerate(bufs):
            ctx.map_split_ptr(i, t.data_ptr(), t.numel())

    # first job: exact-size shuffle (one host sync) that measures the segments; after it either the
    # peer-to-peer windows or fixed NCCL slots, neither with any host synchronisation
    shuffle_slot = [None]
    shuffle_kind = ["nccl all-to-all, exact sizes"]
    biggest_seen = [0]
    shuffle_stats = {}

    def do_shuffle():
        if shuffle_kind[0].startswith("p2p"):
            ctx.shuffle_p2p()
            return
        biggest = parallel.shuffle(ctx, slot_bytes=shuffle_slot[0], stats=shuffle_stats)
        if shuffle_slot[0] is None:
            biggest_seen[0] = biggest
            shuffle_slot[0] = parallel.slot_for(biggest)
            shuffle_kind[0] = f"one nccl all-to-all (fixed {shuffle_slot[0]}-byte slots)"

    def job(resident):
        ctx.job_begin(keep_input=resident)
        if not resident:
            stage_all()
        ctx.map_finish()
        if world > 1:
            do_shuffle()
        ctx.reduce_all()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # untimed: stage the splits once, run one full job
    job(False)
    info = ctx.job_sync()
    assert info.n_tokens == tokens_local, (info.n_tokens, tokens_local)
    owners = "partition mod ranks"
    balanced = world > 1 and not args.mod_owners
    if balanced:
        # equal shares of the output order per rank instead (10 partitions on 8 ranks would leave two
        # ranks with twice the others' keys); one more exact-size job measures the new segments
        parallel.balance_owners(ctx)
        owners = "key ranges (splitters at equal shares of the first job's rows)"
        shuffle_slot[0] = None
        job(True)
        info = ctx.job_sync()
        assert info.n_tokens == tokens_local
    rows_n, key_bytes = int(info.n_keys), int(info.key_bytes)
    tot = torch.tensor([tokens_local, bytes_local, rows_n, key_bytes], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(tot)
    tokens_total, bytes_total, rows_total, keybytes_total = [int(x) for x in tot.cpu()]
    if world > 1 and args.shuffle != "nccl":
        try:
            slot = parallel.p2p_connect(ctx, parallel.p2p_slot_for(biggest_seen[0]))
            shuffle_kind[0] = f"p2p: records stored straight into the owners' windows over NVLink ({slot}-byte slots), flags, no collective"
        except _native.NativeError as e:
            if args.shuffle == "p2p":
                raise
            if rank == 0:
                print(f"[bench] peer-to-peer windows unavailable ({e}); using NCCL", file=sys.stderr)
    # the steady-state path once, untimed, and THAT job is the one compared with the oracle
    job(True)
    info = ctx.job_sync()
    parity = None if args.no_parity else parity_check(env, ctx, union_splits, info, balanced)

    # ---- device-resident steps
    for _ in range(warmup):
        job(True)
    ctx.job_sync()
    sampler = ClockSampler(env.local_rank)
    if rank == 0:
        sampler.start()
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(steps)]
    barrier()
    t_start = torch.cuda.Event(enable_timing=True); t_end = torch.cuda.Event(enable_timing=True)
    t_start.record()
    launches = 0
    for s in range(steps):
        ctx.job_begin(keep_input=True)
        ev[s][0].record()
        ctx.map_finish()
        ev[s][1].record()
        if world > 1:
            do_shuffle()
        ev[s][3].record()
        ctx.reduce_all()
        ev[s][2].record()
        launches += int(ctx.stats().launches)
    t_end.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    info = ctx.job_sync()
    assert info.n_tokens == tokens_local and int(info.n_keys) == rows_n
    flags = torch.tensor([int(info.flags) & 3, int(info.flags) >> 8], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(flags, op=dist.ReduceOp.MAX)
    reduce_path = "msd units (msd_kernels.cuh) on every rank"
    if int(flags[0].item()) & 1:
        reduce_path = f"lsd radix (reduce_kernels.cuh) on at least one rank (msd decline bits {int(flags[1].item())})"
    ms_total = t_start.elapsed_time(t_end)
    map_ms = [ev[s][0].elapsed_time(ev[s][1]) for s in range(steps)]
    red_ms = [ev[s][1].elapsed_time(ev[s][2]) for s in range(steps)]
    shuf_ms = [ev[s][1].elapsed_time(ev[s][3]) for s in range(steps)]
    t = torch.tensor([ms_total, sum(shuf_ms) / len(shuf_ms), sum(map_ms) / len(map_ms)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t[0].item()) / steps
    out = {"ms_step": ms_step, "shuf_ms_max": float(t[1].item()), "map_ms_max": float(t[2].item()),
           "map_ms": sum(map_ms) / len(map_ms), "red_ms": sum(red_ms) / len(red_ms),
           "tokens_total": tokens_total, "bytes_total": bytes_total, "rows_total": rows_total,
           "keybytes_total": keybytes_total, "tokens_local": tokens_local, "bytes_local": bytes_local,
           "rows_n": rows_n, "key_bytes": key_bytes, "launches": launches, "clocks": clocks, "parity": parity,
           "owners": owners, "shuffle_kind": shuffle_kind[0], "reduce_path": reduce_path, "shuffle_stats": shuffle_stats, "e2e": None,
           "nsplits_local": len(splits)}

    # ---- end to end through the public API: host buffers in, rows out, every step
    if want_e2e:
        n_e2e = max(1, min(steps, 10))
        h2d = d2h = 0

        def pinned_buf(n, dtype):
            return torch.empty(max(int(n), 1), dtype=dtype).pin_memory().numpy()[:int(n)]
        # the step's result lands in pinned host buffers (sized from the checked first job)
        res = _native.Rows(pinned_buf(rows_n, torch.int32).view(np.uint32), pinned_buf(rows_n, torch.int64).view(np.uint64),
                           pinned_buf(rows_n + 1, torch.int64).view(np.uint64), pinned_buf(key_bytes, torch.uint8),
                           pinned_buf(R + 1, torch.int64).view(np.uint64))
        job(False); ctx.rows(res)      # warm-up of the path
        barrier()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        w0 = time.perf_counter()
        e0.record()
        for _ in range(n_e2e):
            job(False)
            ctx.rows(res)
            st = ctx.stats()
            h2d, d2h = int(st.h2d_bytes), int(st.d2h_bytes)
        e1.record()
        barrier()
        wall_ms = (time.perf_counter() - w0) * 1e3
        e2e_ms = max(e0.elapsed_time(e1), wall_ms) / n_e2e
        t = torch.tensor([e2e_ms], dtype=torch.float64, device=dev)
        per_rank = [e2e_ms]
        if world > 1:
            gathered = [torch.zeros_like(t) for _ in range(world)]
            dist.all_gather(gathered, t)
            per_rank = [float(g.item()) for g in gathered]
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out["e2e"] = {"value": tokens_total / (float(t.item()) * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                      "d2h_bytes_per_step": d2h, "steps": n_e2e, "ms_per_step": float(t.item()),
                      "ms_per_step_by_rank": [round(x, 3) for x in per_rank],
                      "h2d_GBs_by_rank": [round(h2d / (x * 1e-3) / 1e9, 1) for x in per_rank],
                      "host_placement": getattr(env, "numa", "not bound")}
    ctx.close()
    del bufs
    return out


def run_reference(args, out):
    """The reference's CPU path (oracle staged = job.lua + utils.merge_iterator restated in C; the Lua +
    MongoDB original cannot run in this image), on all host cores, on the SAME workload as our arm at
    N = 1: every step maps and reduces all 197 splits of config 2."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from lua_mapreduce_1_b200 import corpus
    from oracle import oracle as O
    threads = host_threads()
    lines = corpus.split_lines(args.scale)
    splits, tokens = corpus.generate(lines=lines, seed=SEED)
    for _ in range(args.warmup):
        O.run_staged(splits, R, threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        r = O.run_staged(splits, R, threads)
        assert r.tokens == tokens
    dt = (time.perf_counter() - t0) / max(args.steps, 1)
    value = tokens / dt
    sample_txt = (f"all {len(splits)} splits of config 2 ({tokens} tokens) per step, {args.steps} steps, oracle staged path "
                  f"(job.lua + merge_iterator restated in C), {threads} OpenMP threads as workers")
    out.write(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8/u64", "data": "synthetic",
        "config": {"workload": f"config2 x{args.scale}: synthetic Europarl-shaped word count, {len(splits)} splits/rank, R={R}",
                   "tokens": tokens, "partitions": R, "seed": hex(SEED),
                   "note": "the Lua+MongoDB reference cannot run in this image (no Lua VM, no mongod); "
                           "this is its CPU algorithm restated in C with OpenMP threads standing in for workers; "
                           "one config-2 shard is timed at every N (the host cores' tokens/s does not depend on how "
                           "many shards they are given)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample_txt},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }) + "\n")


def main():
    args = parse()
    # stdout carries ONE JSON line: whatever libraries print there while the job runs (NCCL's version
    # banner, for one) goes to stderr; the line itself is written to the saved descriptor at the end
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    try:
        if args.impl == "reference":
            return run_reference(args, real_stdout)
        return run_ours(args, real_stdout)
    finally:
        real_stdout.flush()


def run_ours(args, out):

    import torch
    import torch.distributed as dist
    from lua_mapreduce_1_b200 import corpus

    env = Env()
    env.world = int(os.environ.get("WORLD_SIZE", "1"))
    env.rank = int(os.environ.get("RANK", "0"))
    env.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world, rank = env.world, env.rank
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the data plane has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(env.local_rank)
    env.dev = torch.device("cuda", env.local_rank)
    env.numa = bind_to_gpu_numa(env.local_rank) if world > 1 and not os.environ.get("MRGPU_NO_NUMA_BIND") else "not bound"
    if world > 1:
        dist.init_process_group("nccl", device_id=env.dev)
    assert world == args.gpus or world == 1, "launch with torchrun --nproc-per-node N for --gpus N"
    # a non-default stream: the library treats a NULL handle as "create a private stream", and the
    # events below must be recorded on the stream the kernels run on
    env.stream = torch.cuda.Stream(device=env.dev)
    torch.cuda.set_stream(env.stream)

    # ---- headline: config 2, one 197-split shard per rank (split ids are global, bytes depend only on the id)
    lines = corpus.split_lines(args.scale)
    nper = len(lines)
    splits, tokens_local = corpus.generate(lines=lines, seed=SEED, first_split=rank * nper)
    m = run_workload(env, args, splits, tokens_local,
                     lambda: corpus.generate(lines=lines * world, seed=SEED)[0] if world > 1 else splits,
                     args.steps, args.warmup, not args.no_e2e, pin=True)

    # ---- strong scaling: the fixed config-3 corpus, splits dealt round-robin to the ranks
    strong = None
    if not args.no_strong:
        lines3 = corpus.split_lines(CONFIG3_SCALE)
        ids = list(range(rank, len(lines3), world))
        steps3, warm3 = max(3, min(args.steps, 10)), max(3, min(args.warmup, 3))
        sp3, tok3 = [], 0
        for i in ids:
            s, tk = corpus.generate_split(i, lines3[i], SEED)
            sp3.append(s); tok3 += tk
        del splits
        s3 = run_workload(env, args, sp3, tok3, lambda: corpus.generate(lines=lines3, seed=SEED)[0] if world > 1 else sp3,
                          steps3, warm3, False, pin=False)
        strong = {"workload": f"config3: fixed corpus of {len(lines3)} splits (config 2 x{CONFIG3_SCALE}), split i on rank i mod {world}",
                  "scaling": "strong", "tokens": s3["tokens_total"], "input_bytes": s3["bytes_total"],
                  "distinct_keys": s3["rows_total"], "splits_on_rank0": s3["nsplits_local"],
                  "ms_per_step": s3["ms_step"], "value": s3["tokens_total"] / (s3["ms_step"] * 1e-3), "unit": UNIT,
                  "steps": steps3, "warmup": warm3, "map_ms_max": s3["map_ms_max"],
                  "exchange_and_reduce_ms_rank0": s3["red_ms"], "reduce_path": s3["reduce_path"],
                  "efficiency_vs_n1_ms": "t_1 / (N * ms_per_step), t_1 = this field of the N=1 line",
                  "parity_check": s3["parity"]}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (map_combine_kernel), SURVEY.md 8d arithmetic
    peak, peak_src = peaks()
    alg_bytes = m["bytes_local"] + m["key_bytes"] + 8 * m["rows_n"]          # per launch on this rank: B_in + B_out
    map_avg_ms = m["map_ms"]
    achieved = alg_bytes / (map_avg_ms * 1e-3) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(tp):
        try:
            traffic = json.load(open(tp)).get("map_combine_kernel_dram_bytes_per_launch")
        except Exception:
            traffic = None
    ms_step = m["ms_step"]
    roofline = {"bound": "hbm", "kernel": "map_combine_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg_bytes, "kernel_ms": map_avg_ms,
                "job_frac": (m["bytes_total"] + m["keybytes_total"] + 8 * m["rows_total"]) / world / (ms_step * 1e-3) / 1e9 / peak,
                "map_ms": map_avg_ms, "reduce_ms": m["red_ms"]}

    # the exchange against the NVLink roofline: bytes that leave this rank per job over
    # the time from the end of its map side to the end of the exchange (waiting for the slowest
    # sender included), max over ranks; 900 GB/s per direction per GPU
    nvlink = None
    if world > 1 and "matrix" in m["shuffle_stats"]:
        mtx = m["shuffle_stats"]["matrix"][rank]
        away = [d for d in range(world) if d != rank]
        nrec_away = int(sum(int(mtx[d]) for d in away))
        long_away = int(sum(int(mtx[world + d]) for d in away))
        mean_key = m["keybytes_total"] / max(m["rows_total"], 1)
        rec_bytes = 32 if m["shuffle_kind"].startswith("p2p") else 24
        shuf = m["shuf_ms_max"]
        nvlink = {"shuffle_ms": shuf, "records_leaving_rank0": nrec_away,
                  "wire_bytes_rank0": rec_bytes * nrec_away + long_away,
                  "algorithmic_bytes_rank0": int(nrec_away * (mean_key + 8)),
                  "achieved": nrec_away * (mean_key + 8) / (shuf * 1e-3) / 1e9, "peak": 900.0, "unit": "GB/s",
                  "note": "algorithmic = (mean key length of the job + 8) per record leaving the rank"}
        nvlink["frac"] = nvlink["achieved"] / nvlink["peak"]
        nvlink["wire_frac"] = nvlink["wire_bytes_rank0"] / (shuf * 1e-3) / 1e9 / 900.0

    cpu = None
    if not args.no_cpu_baseline and world == 1:
        from oracle import oracle as O
        threads = host_threads()
        splits, _ = corpus.generate(lines=lines, seed=SEED)
        t0 = time.perf_counter()
        r = O.run_staged(splits, R, threads)
        dt = time.perf_counter() - t0
        assert r.tokens == m["tokens_local"]
        cpu = {"value": m["tokens_local"] / dt, "unit": UNIT, "cores": threads, "kind": "port",
               "sample": f"all {len(splits)} splits of config 2 once ({m['tokens_local']} tokens, {dt:.2f} s), "
                         "oracle staged path with OpenMP threads as workers"}

    out.write(json.dumps({
        "metric": METRIC, "value": m["tokens_total"] / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8/u64", "data": "synthetic",
        "config": {"workload": f"config2 x{args.scale}: synthetic Europarl-shaped word count, {nper} splits/rank, R={R}",
                   "tokens": m["tokens_total"], "input_bytes": m["bytes_total"], "distinct_keys": m["rows_total"],
                   "partitions": R, "seed": hex(SEED), "parallelism": f"splits sharded over {world} rank(s), "
                   f"owner = {m['owners']}" + (f", shuffle = {m['shuffle_kind']}" if world > 1 else ""),
                   "reduce_path": m["reduce_path"],
                   "l2": "input per rank (294 MB) is larger than L2 (126 MB); no flush between steps"},
        "roofline": roofline, "nvlink": nvlink, "cpu_baseline": cpu, "e2e": m["e2e"], "gpu_launches": m["launches"],
        "clocks": m["clocks"], "parity_check": m["parity"], "strong_config3": strong,
    }) + "\n")
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
