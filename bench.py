#!/usr/bin/env python
"""bench.py -- word-count tokens/s (map + shuffle + reduce) on B200, BASELINE.json's metric.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

Workload: BASELINE.json configs[1], the synthetic Europarl-shaped corpus (197 splits x
10,000 lines, ~49 M Zipf tokens, R = 10).  With N ranks every rank maps its own
197-split shard (weak scaling) and the partitions are merged through one all-to-all.
A step = one whole job over the resident splits: clear tables, tokenise + combine,
partition, (shuffle,) radix sort, segmented reduce, rows left in HBM.
Prints ONE JSON line (rank 0).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "wordcount_tokens_per_s"
UNIT = "tokens/s"
R = 10
SEED = 0x5EED0001


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
    return ap.parse_args()


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


def cpu_reference(splits, tokens, threads, repeats=1):
    """The reference's CPU path restated (oracle/, staged = job.lua + utils.merge_iterator), timed."""
    from oracle import oracle as O
    best = None
    for _ in range(repeats):
        t0 = time.perf_counter()
        r = O.run_staged(splits, R, threads)
        dt = time.perf_counter() - t0
        assert r.tokens == tokens
        best = dt if best is None else min(best, dt)
    return tokens / best, best


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from lua_mapreduce_1_b200 import corpus
    from oracle import oracle as O
    threads = O.max_threads()
    lines = corpus.split_lines(1)
    sample = lines if args.steps + args.warmup <= 12 else lines[:48]
    splits, tokens = corpus.generate(lines=sample, seed=SEED)
    for _ in range(args.warmup):
        O.run_staged(splits, R, threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        r = O.run_staged(splits, R, threads)
    dt = (time.perf_counter() - t0) / max(args.steps, 1)
    value = tokens / dt
    sample_txt = f"{len(sample)} of 197 splits ({tokens} tokens) per step, oracle staged path (job.lua + merge_iterator restated in C)"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8/u64", "data": "synthetic",
        "config": {"workload": "config2: synthetic Europarl-shaped word count, 197 splits x 10k lines, R=10",
                   "note": "the Lua+MongoDB reference cannot run in this image (no Lua VM, no mongod); "
                           "this is its CPU algorithm restated in C with OpenMP threads standing in for workers; "
                           "one config-2 shard is timed at every N (the host cores' tokens/s does not depend on how "
                           "many shards they are given)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample_txt},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)

    import numpy as np
    import torch
    import torch.distributed as dist
    from lua_mapreduce_1_b200 import _native, corpus, parallel

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the data plane has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    assert world == args.gpus or world == 1, "launch with torchrun --nproc-per-node N for --gpus N"

    # ---- corpus: every rank generates its own shard (split ids are global, bytes depend only on the id)
    lines = corpus.split_lines(args.scale)
    nper = len(lines)
    splits, tokens_local = corpus.generate(lines=lines, seed=SEED, first_split=rank * nper)
    bytes_local = sum(len(s) for s in splits)
    # a non-default stream: the library treats a NULL handle as "create a private stream", and the
    # events below must be recorded on the stream the kernels run on
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    max_keys = 1 << 20
    ctx = _native.Context(R, bytes_local + 64 * len(splits) + (1 << 16), max_keys=max_keys * (2 if world > 1 else 1),
                          table_slots=1 << 21, device=local_rank, stream=stream.cuda_stream, world_size=world, rank=rank)
    pinned = [torch.from_numpy(s).pin_memory() for s in splits]

    def stage_all():
        for i, t in enumerate(pinned):
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

    # untimed: stage the splits once, run one full job and check it
    job(False)
    info = ctx.job_sync()
    assert info.n_tokens == tokens_local, (info.n_tokens, tokens_local)
    owners = "partition mod ranks"
    if world > 1 and not args.mod_owners:
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

    # ---- device-resident steps
    for _ in range(args.warmup):
        job(True)
    ctx.job_sync()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(args.steps)]
    barrier()
    t_start = torch.cuda.Event(enable_timing=True); t_end = torch.cuda.Event(enable_timing=True)
    t_start.record()
    launches = 0
    for s in range(args.steps):
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
    assert info.n_tokens == tokens_local
    ms_total = t_start.elapsed_time(t_end)
    map_ms = [ev[s][0].elapsed_time(ev[s][1]) for s in range(args.steps)]
    red_ms = [ev[s][1].elapsed_time(ev[s][2]) for s in range(args.steps)]
    shuf_ms = [ev[s][1].elapsed_time(ev[s][3]) for s in range(args.steps)]
    t = torch.tensor([ms_total, sum(shuf_ms) / len(shuf_ms)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t[0].item()) / args.steps
    shuf_ms_max = float(t[1].item())
    value = tokens_total / (ms_step * 1e-3)

    # ---- end to end through the public API: host buffers in, rows out, every step
    e2e = None
    if not args.no_e2e:
        n_e2e = max(1, min(args.steps, 10))
        h2d = d2h = 0
        def pinned_buf(n, dtype):
            return torch.empty(max(int(n), 1), dtype=dtype).pin_memory().numpy()[:int(n)]
        # the step's result lands in pinned host buffers (sized from the checked first job)
        out = _native.Rows(pinned_buf(rows_n, torch.int32).view(np.uint32), pinned_buf(rows_n, torch.int64).view(np.uint64),
                           pinned_buf(rows_n + 1, torch.int64).view(np.uint64), pinned_buf(key_bytes, torch.uint8),
                           pinned_buf(R + 1, torch.int64).view(np.uint64))
        job(False); ctx.rows(out)      # warm-up of the path
        barrier()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        w0 = time.perf_counter()
        e0.record()
        for _ in range(n_e2e):
            job(False)
            ctx.rows(out)
            st = ctx.stats()
            h2d, d2h = int(st.h2d_bytes), int(st.d2h_bytes)
        e1.record()
        barrier()
        wall_ms = (time.perf_counter() - w0) * 1e3
        e2e_ms = max(e0.elapsed_time(e1), wall_ms) / n_e2e
        t = torch.tensor([e2e_ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e = {"value": tokens_total / (float(t.item()) * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h, "steps": n_e2e, "ms_per_step": float(t.item())}

    if rank != 0:
        ctx.close()
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (map_combine_kernel), SURVEY.md 8d arithmetic
    peak, peak_src = peaks()
    alg_bytes = bytes_local + key_bytes + 8 * rows_n          # per launch on this rank: B_in + B_out
    map_avg_ms = sum(map_ms) / len(map_ms)
    achieved = alg_bytes / (map_avg_ms * 1e-3) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(tp):
        try:
            traffic = json.load(open(tp)).get("map_combine_kernel_dram_bytes_per_launch")
        except Exception:
            traffic = None
    roofline = {"bound": "hbm", "kernel": "map_combine_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg_bytes, "kernel_ms": map_avg_ms,
                "job_frac": (bytes_total + keybytes_total + 8 * rows_total) / world / (ms_step * 1e-3) / 1e9 / peak,
                "map_ms": map_avg_ms, "reduce_ms": sum(red_ms) / len(red_ms)}

    # the exchange against the NVLink roofline: bytes that leave this rank per job over
    # the time from the end of its map side to the end of the exchange (waiting for the slowest
    # sender included), max over ranks; 900 GB/s per direction per GPU
    nvlink = None
    if world > 1 and "matrix" in shuffle_stats:
        mtx = shuffle_stats["matrix"][rank]
        away = [d for d in range(world) if d != rank]
        nrec_away = int(sum(int(mtx[d]) for d in away))
        long_away = int(sum(int(mtx[world + d]) for d in away))
        mean_key = keybytes_total / max(rows_total, 1)
        rec_bytes = 32 if shuffle_kind[0].startswith("p2p") else 24
        nvlink = {"shuffle_ms": shuf_ms_max, "records_leaving_rank0": nrec_away,
                  "wire_bytes_rank0": rec_bytes * nrec_away + long_away,
                  "algorithmic_bytes_rank0": int(nrec_away * (mean_key + 8)),
                  "achieved": nrec_away * (mean_key + 8) / (shuf_ms_max * 1e-3) / 1e9, "peak": 900.0, "unit": "GB/s",
                  "note": "algorithmic = (mean key length of the job + 8) per record leaving the rank"}
        nvlink["frac"] = nvlink["achieved"] / nvlink["peak"]
        nvlink["wire_frac"] = nvlink["wire_bytes_rank0"] / (shuf_ms_max * 1e-3) / 1e9 / 900.0

    cpu = None
    if not args.no_cpu_baseline and world == 1:
        from oracle import oracle as O
        threads = O.max_threads()
        v, dt = cpu_reference(splits, tokens_local, threads)
        cpu = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
               "sample": f"all {len(splits)} splits of config 2 once ({tokens_local} tokens, {dt:.2f} s), "
                         "oracle staged path with OpenMP threads as workers"}

    print(json.dumps({
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8/u64", "data": "synthetic",
        "config": {"workload": f"config2 x{args.scale}: synthetic Europarl-shaped word count, {nper} splits/rank, R={R}",
                   "tokens": tokens_total, "input_bytes": bytes_total, "distinct_keys": rows_total,
                   "partitions": R, "seed": hex(SEED), "parallelism": f"splits sharded over {world} rank(s), "
                   f"owner = {owners}" + (f", shuffle = {shuffle_kind[0]}" if world > 1 else ""),
                   "l2": "input per rank (294 MB) is larger than L2 (126 MB); no flush between steps"},
        "roofline": roofline, "nvlink": nvlink, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
    }))
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
