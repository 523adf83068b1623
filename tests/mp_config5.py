"""BASELINE.json configs[4] as it is stated: skewed integer-key group-by-sum, 10^9 records over ALL ranks
(one rank per GPU, torchrun), heavy-hitter keys, uneven partitions.  Every rank draws its share of the
records (the same generator as tests/run_config.py --config 5, a different stream per rank), maps them
with mr_map_records, the combined (key, sum) records cross through the peer-to-peer windows (NCCL
where the GPUs cannot map each other), every owner reduces.  Checked against the oracle's group-by
over the union of all records (rank 0, all host threads): an order-independent digest of every
(partition, key, sum) row, the row count, the sum of sums; rows on every rank in numeric key order.
Prints one JSON line (rank 0).

    torchrun --nproc-per-node 8 tests/mp_config5.py [--records 1e9]
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # noqa: E402  (rows_digest64)
from lua_mapreduce_1_b200 import _native, parallel  # noqa: E402

R = 10


def records_of_rank(rank, n):
    rng = np.random.default_rng([0x5EED0005, rank])
    keys = (rng.zipf(1.2, n) % (1 << 22)).astype(np.uint32) * np.uint32(2654435761)
    hot = rng.random(n)
    for j, k in enumerate((7, 0, 0xFFFFFFFF, 123456789)):
        keys[(hot >= 0.06 * j) & (hot < 0.06 * (j + 1))] = k          # four heavy hitters, 6 % of the records each
    vals = rng.integers(0, 256, n, dtype=np.uint32)
    return keys, vals


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--records", type=float, default=1e9)
    ap.add_argument("--steps", type=int, default=5)
    a = ap.parse_args()
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    rank = dist.get_rank() if world > 1 else 0
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    n = int(a.records) // world
    keys, vals = records_of_rank(rank, n)
    ctx = _native.Context(R, 8 * n + (1 << 20), max_keys=1 << 23, table_slots=1 << 24, device=local_rank,
                          stream=stream.cuda_stream, world_size=world, rank=rank, map_kind=_native.MR_MAP_RECORDS_U32,
                          partition_kind=_native.MR_PARTITION_MOD_U32)
    batches = list(zip(np.array_split(keys, 16), np.array_split(vals, 16)))
    kind = ["single rank"]
    state = {"slot": None, "p2p": False}

    def shuffle():
        if world == 1:
            return
        if state["p2p"]:
            ctx.shuffle_p2p()
            return
        biggest = parallel.shuffle(ctx, slot_bytes=state["slot"])
        if state["slot"] is None:
            state["biggest"] = biggest

    def job(first):
        if first:
            for i, (k, v) in enumerate(batches):
                ctx.map_records(i, k, v)
        else:
            ctx.job_begin(keep_input=True)
        ctx.map_finish()
        shuffle()
        ctx.reduce_all()

    job(True)                                    # exact-size shuffle: measures the segments
    info = ctx.job_sync()
    if world > 1:
        try:
            parallel.p2p_connect(ctx, parallel.p2p_slot_for(state["biggest"]))
            state["p2p"] = True
            kind[0] = "p2p windows"
        except _native.NativeError:
            state["slot"] = parallel.slot_for(state["biggest"])
            kind[0] = "nccl fixed slots"
        job(False)
        info = ctx.job_sync()
    assert int(info.n_tokens) == n
    rows = ctx.rows()
    d, nrows, ssum = bench.rows_digest64(rows.part, rows.count, rows.key_off, rows.key_bytes)
    ku = rows.keys_u32().astype(np.int64)
    part = rows.part.astype(np.int64)
    ordered = bool(np.all((np.diff(part) > 0) | ((np.diff(part) == 0) & (np.diff(ku) > 0)))) and bool(np.all(ku % R == part))
    mine = torch.from_numpy(np.array([d, nrows, ssum, int(ordered)], dtype=np.uint64).view(np.int64).copy()).to(dev)
    if world > 1:
        dist.all_reduce(mine)
    got = [int(v) & 0xFFFFFFFFFFFFFFFF for v in mine.cpu().numpy().view(np.uint64)]
    # timed resident jobs
    for _ in range(2):
        job(False)
    ctx.job_sync()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        job(False)
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / a.steps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    flags = int(ctx.job_sync().flags)
    ctx.close()
    ok = True
    if rank == 0:
        from oracle import oracle as O
        t0 = time.perf_counter()
        allrec = [records_of_rank(r, n) for r in range(world)]
        want = O.groupby_u32(allrec, R)
        wd, wn, wsum = bench.rows_digest64(want.part, want.count, want.key_off, want.key_bytes)
        ok = got[0] == wd and got[1] == wn and got[2] == wsum and got[3] == world
        print(json.dumps({"config": 5, "workload": "integer-key group-by-sum, Zipf(1.2) keys + 4 heavy hitters (6 % each), R=10",
                          "n_gpus": world, "records": n * world, "distinct_keys": got[1], "bit_exact": bool(ok),
                          "digest": f"{got[0]:016x}", "oracle_digest": f"{wd:016x}", "sum_of_sums": got[2],
                          "ranks_in_order": got[3], "shuffle": kind[0], "resident_job_ms": ms,
                          "records_per_s": n * world / (ms * 1e-3), "per_gpu_input_GBs": 8 * n / (ms * 1e-3) / 1e9,
                          "reduce_path_flags_rank0": flags, "oracle_s": round(time.perf_counter() - t0, 1)}))
    if world > 1:
        v = [ok]
        dist.broadcast_object_list(v, src=0)
        ok = v[0]
        dist.destroy_process_group()
    if not ok:
        sys.exit(1)


if __name__ == "__main__":
    main()
