"""Run under torchrun, one rank per GPU: randomised cases (tests/test_gpu_fuzz.make_case) through the
multi-GPU paths -- peer-to-peer shuffle and NCCL fixed slots, partition-mod-world owners and balanced
key-range owners -- each rank's rows against its piece of the whole-job oracle, bit-exact."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from lua_mapreduce_1_b200 import _native, parallel  # noqa: E402
from oracle import oracle as O  # noqa: E402
from test_gpu_fuzz import make_case, total_bytes  # noqa: E402


def check(ctx, want, owner, what):
    got = ctx.rows()
    mine = owner == ctx.rank
    assert got.n == int(mine.sum()), (what, ctx.rank, got.n, int(mine.sum()))
    assert got.keys() == [k for k, o in zip(want.keys(), mine) if o], f"{what}: rank {ctx.rank}: keys differ"
    assert np.array_equal(got.part, want.part[mine]) and np.array_equal(got.count, want.count[mine]), what


def main():
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist.init_process_group("nccl", device_id=dev)
    rank, world = dist.get_rank(), dist.get_world_size()
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    ncases = int(os.environ.get("MRGPU_FUZZ_CASES", "24"))
    for seed in range(300, 300 + ncases):
        splits, R = make_case(seed)
        want = O.run_naive(splits, R)
        ctx = _native.Context(R, total_bytes(splits) + 64 * len(splits) + 4096, max_keys=1 << 16, device=local_rank,
                              stream=stream.cuda_stream, world_size=world, rank=rank)
        for i in parallel.shard_splits(len(splits), world, rank):
            ctx.map_split(i, splits[i])
        ctx.map_finish()
        biggest = parallel.shuffle(ctx)
        ctx.reduce_all()
        owner = np.array([int(p) % world for p in want.part], dtype=np.int64)
        check(ctx, want, owner, f"seed {seed} exact")
        parallel.p2p_connect(ctx, parallel.p2p_slot_for(max(biggest, 64)))
        for mode in ("p2p", "p2p", "fixed"):
            ctx.job_begin(keep_input=True)
            ctx.map_finish()
            if mode == "p2p":
                ctx.shuffle_p2p()
            else:
                parallel.shuffle(ctx, slot_bytes=parallel.slot_for(max(biggest, 64)))
            ctx.reduce_all()
            check(ctx, want, owner, f"seed {seed} {mode}")
        if want.n:
            parallel.balance_owners(ctx)
            owner = np.array([ctx.owner_of(k) for k in want.keys()], dtype=np.int64)
            assert np.all(np.diff(owner) >= 0), seed
            ctx.job_begin(keep_input=True)
            ctx.map_finish()
            biggest = parallel.shuffle(ctx)                   # exact sizes under the new owners
            ctx.reduce_all()
            check(ctx, want, owner, f"seed {seed} balanced exact")
            parallel.p2p_connect(ctx, parallel.p2p_slot_for(max(biggest, 64)))
            for _ in range(2):
                ctx.job_begin(keep_input=True)
                ctx.map_finish()
                ctx.shuffle_p2p()
                ctx.reduce_all()
                check(ctx, want, owner, f"seed {seed} balanced p2p")
        ctx.job_sync()
        dist.barrier()
        ctx.close()
        dist.barrier()
    dist.destroy_process_group()
    print(f"rank {rank}/{world}: multi-GPU fuzz ok ({ncases} cases)")


if __name__ == "__main__":
    main()
