"""Run under torchrun with one rank per GPU (nccl): multi-GPU parity of the whole job with the oracle.
Every rank maps its shard, the partitions are exchanged with ONE all-to-all, the owner merges; the
rows of every owned partition must equal the whole-job oracle rows of that partition, in order."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from lua_mapreduce_1_b200 import _native, corpus, parallel  # noqa: E402
from oracle import oracle as O  # noqa: E402


def main():
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist.init_process_group("nccl", device_id=dev)
    rank, world = dist.get_rank(), dist.get_world_size()
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    for R, lines in ((10, [2000, 1500, 1, 0, 1800, 900, 2100, 333, 1200]), (15, [500] * 5)):
        splits, _ = corpus.generate(lines=lines, threads=4)
        splits.append(np.frombuffer(corpus.TORTURE, dtype=np.uint8))
        want = O.run_staged(splits, R, 4)
        ctx = _native.Context(R, 1 << 24, max_keys=1 << 19, device=local_rank, stream=stream.cuda_stream,
                              world_size=world, rank=rank)
        slot = None
        for it in range(3):                      # later passes re-use the resident splits (keep_input)
            if it == 0:
                for i in parallel.shard_splits(len(splits), world, rank):
                    ctx.map_split(i, splits[i])
            else:
                ctx.job_begin(keep_input=True)
            ctx.map_finish()
            if it < 2:                           # exact sizes first, then fixed slots (no host sync)
                biggest = parallel.shuffle(ctx, slot_bytes=slot)
                slot = parallel.slot_for(biggest)
            else:                                # a slot that cannot hold the segments must be reported
                parallel.shuffle(ctx, slot_bytes=64)
                ctx.reduce_all()
                try:
                    ctx.job_sync()
                    raise AssertionError("overflowing shuffle slot was not reported")
                except _native.NativeError as e:
                    assert e.code == _native.MR_ERR_OVERFLOW, e
                continue
            ctx.reduce_all()
            got = ctx.rows()
            owned = np.array([parallel.owner_of_partition(int(p), world) == rank for p in want.part], dtype=bool)
            keys = want.keys()
            wk = [k for k, o in zip(keys, owned) if o]
            assert got.n == int(owned.sum()), (rank, got.n, int(owned.sum()))
            assert got.keys() == wk, f"rank {rank}: keys / order differ"
            assert np.array_equal(got.part, want.part[owned]) and np.array_equal(got.count, want.count[owned])
            tot = torch.tensor([int(got.count.sum())], dtype=torch.int64, device=dev)
            dist.all_reduce(tot)
            assert int(tot.item()) == want.tokens
        ctx.close()
    # number keys with a MIN reducer: the owner really merges (the same key arrives from several ranks)
    rng = np.random.default_rng(17)
    batches = [(rng.integers(0, 20000, 50000).astype(np.uint32), rng.integers(0, 2**32, 50000, dtype=np.uint64).astype(np.uint32))
               for _ in range(2 * world + 1)]
    for kind, op in ((_native.MR_REDUCE_MIN_U64, 1), (_native.MR_REDUCE_SUM_U64, 0)):
        want = O.groupby_u32(batches, 10, op)
        ctx = _native.Context(10, 1 << 24, max_keys=1 << 18, device=local_rank, stream=stream.cuda_stream,
                              world_size=world, rank=rank, map_kind=_native.MR_MAP_RECORDS_U32,
                              partition_kind=_native.MR_PARTITION_MOD_U32, reduce_kind=kind)
        for i in parallel.shard_splits(len(batches), world, rank):
            ctx.map_records(i, batches[i][0], batches[i][1])
        ctx.map_finish()
        parallel.shuffle(ctx)
        ctx.reduce_all()
        got = ctx.rows()
        owned = np.array([parallel.owner_of_partition(int(p), world) == rank for p in want.part], dtype=bool)
        assert got.n == int(owned.sum()) and np.array_equal(got.count, want.count[owned]), (rank, kind)
        assert np.array_equal(got.keys_u32(), np.frombuffer(want.key_bytes.tobytes(), ">u4")[owned])
        ctx.close()
    # the library's own transport (mr_nccl_init / mr_shuffle_nccl): torch only carries the rendezvous
    # blob here, as the host's control plane would; the all-to-all runs inside libmrgpu.so
    blob = [_native.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(blob, src=0)
    splits, _ = corpus.generate(lines=[1500, 1700, 1, 0, 900, 2500, 640], threads=4)
    splits.append(np.frombuffer(corpus.TORTURE, dtype=np.uint8))
    want = O.run_staged(splits, 10, 4)
    ctx = _native.Context(10, 1 << 24, max_keys=1 << 19, device=local_rank, stream=stream.cuda_stream,
                          world_size=world, rank=rank)
    ctx.nccl_init(blob[0])
    slot = 0
    for it in range(3):
        if it == 0:
            for i in parallel.shard_splits(len(splits), world, rank):
                ctx.map_split(i, splits[i])
        else:
            ctx.job_begin(keep_input=True)
        ctx.map_finish()
        used = ctx.shuffle_nccl(slot if it < 2 else 64)   # sized from the job, then re-used, then too small
        ctx.reduce_all()
        if it == 2:
            try:
                ctx.job_sync()
                raise AssertionError("overflowing slot was not reported by the native transport")
            except _native.NativeError as e:
                assert e.code == _native.MR_ERR_OVERFLOW, e
            continue
        assert used >= 4096 and used % 4096 == 0 and (it == 0 or used == slot), (it, used, slot)
        slot = used
        got = ctx.rows()
        owned = np.array([parallel.owner_of_partition(int(p), world) == rank for p in want.part], dtype=bool)
        assert got.n == int(owned.sum()), (rank, got.n, int(owned.sum()))
        assert got.keys() == [k for k, o in zip(want.keys(), owned) if o], f"rank {rank}: native transport keys differ"
        assert np.array_equal(got.part, want.part[owned]) and np.array_equal(got.count, want.count[owned])
    ctx.close()
    # peer-to-peer shuffle: records stored straight into the owners' windows (CUDA IPC), flags, no
    # collective call in the data path; five jobs so that both buffers of the window are re-used
    ctx = _native.Context(10, 1 << 24, max_keys=1 << 19, device=local_rank, stream=stream.cuda_stream,
                          world_size=world, rank=rank)
    for i in parallel.shard_splits(len(splits), world, rank):
        ctx.map_split(i, splits[i])
    ctx.map_finish()
    biggest = parallel.shuffle(ctx)                       # exact sizes once, to size the slots
    ctx.reduce_all(); ctx.job_sync()
    parallel.p2p_connect(ctx, parallel.p2p_slot_for(biggest))
    owned = np.array([parallel.owner_of_partition(int(p), world) == rank for p in want.part], dtype=bool)
    wk = [k for k, o in zip(want.keys(), owned) if o]
    for it in range(5):
        ctx.job_begin(keep_input=True)
        ctx.map_finish()
        ctx.shuffle_p2p()
        ctx.reduce_all()
        got = ctx.rows()
        assert got.n == int(owned.sum()), (rank, it, got.n, int(owned.sum()))
        assert got.keys() == wk, f"rank {rank}: p2p keys differ in job {it}"
        assert np.array_equal(got.part, want.part[owned]) and np.array_equal(got.count, want.count[owned])
    # owners by key range: splitters at equal shares of this job's rows; every later job, through any
    # of the three shuffles, leaves rank r with the r-th contiguous piece of the oracle's output order
    splitters = parallel.balance_owners(ctx)
    assert len(splitters) == world - 1 and splitters == sorted(splitters)
    owner = np.array([ctx.owner_of(k) for k in want.keys()])
    assert np.all(np.diff(owner) >= 0), "owner ranks must follow the output order"
    share = np.bincount(owner, minlength=world)
    assert share.max() - share.min() <= max(8, want.n // 50), share      # equal shares (keys equal in 7 bytes stay together)
    mine = owner == rank
    wk = [k for k, o in zip(want.keys(), mine) if o]
    slot = None
    for mode in ("exact", "fixed", "p2p", "p2p"):
        ctx.job_begin(keep_input=True)
        ctx.map_finish()
        if mode == "p2p":
            ctx.shuffle_p2p()
        else:
            biggest = parallel.shuffle(ctx, slot_bytes=slot)
            slot = parallel.slot_for(biggest)
        ctx.reduce_all()
        got = ctx.rows()
        assert got.n == int(mine.sum()), (rank, mode, got.n, int(mine.sum()))
        assert got.keys() == wk, f"rank {rank}: keys differ with range owners, {mode} shuffle"
        assert np.array_equal(got.part, want.part[mine]) and np.array_equal(got.count, want.count[mine])
        if mode == "fixed":
            parallel.p2p_connect(ctx, parallel.p2p_slot_for(biggest))
    ctx.job_sync()
    ctx.set_owners([])                                    # back to partition mod world
    dist.barrier()
    parallel.p2p_connect(ctx, 64)                         # slots that cannot hold the segments
    ctx.job_begin(keep_input=True)
    ctx.map_finish()
    ctx.shuffle_p2p()
    ctx.reduce_all()
    try:
        ctx.job_sync()
        raise AssertionError("overflowing p2p slot was not reported")
    except _native.NativeError as e:
        assert e.code == _native.MR_ERR_OVERFLOW, e
    dist.barrier()
    ctx.close()
    dist.barrier()
    dist.destroy_process_group()
    print(f"rank {rank}/{world}: multi-GPU parity ok")


if __name__ == "__main__":
    main()
