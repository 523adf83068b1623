"""Run under torchrun (gloo, CPU): the host-side shuffle logic of parallel.py with world_size ranks.

Each rank plays one GPU rank with the ORACLE standing in for the device kernels (this is test code):
it counts its shard of the splits, packs (key, count) records per owner rank in the library's wire
format, moves them with parallel.exchange_plan / exchange_segments, merges what it received, and
checks its partitions against the whole-job oracle."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from lua_mapreduce_1_b200 import corpus, parallel  # noqa: E402
from oracle import oracle as O  # noqa: E402
from wire import pack_segments, unpack_segments  # noqa: E402

R = 10


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    lines = [300, 1, 250, 0, 400, 123, 77]
    splits, _ = corpus.generate(lines=lines, threads=2)
    splits.append(np.frombuffer(corpus.TORTURE, dtype=np.uint8))      # long keys travel as bytes
    mine = parallel.shard_splits(len(splits), world, rank)
    assert sorted(sum((parallel.shard_splits(len(splits), world, r) for r in range(world)), [])) == \
        list(range(len(splits)))
    local = O.run_naive([splits[i] for i in mine], R)                 # map + combine of this rank
    send, send_off, send_nrec, send_lbytes = pack_segments(local, world)
    counts = torch.from_numpy(np.concatenate([send_nrec, send_lbytes]).astype(np.int64))
    matrix = parallel.gather_counts(counts)
    plan_send_off, recv_off, recv_nrec = parallel.plan_from_counts(matrix, rank)
    assert np.array_equal(plan_send_off, send_off)        # the plan predicts the packed layout
    recv = torch.empty(max(int(recv_off[-1]), 1), dtype=torch.uint8)
    parallel.exchange_segments(torch.from_numpy(send), send_off, recv, recv_off)
    merged = unpack_segments(recv.numpy(), recv_off, recv_nrec)       # {key: count}, summed over sources
    want = O.run_naive(splits, R)                                      # whole job
    owned = {k: c for p, k, c in zip(want.part, want.keys(), (int(c) for c in want.count))
             if parallel.owner_of_partition(int(p), world) == rank}
    assert merged == owned, (len(merged), len(owned))
    # every partition has exactly one owner and nothing is lost
    tot = torch.tensor([sum(merged.values())], dtype=torch.int64)
    dist.all_reduce(tot)
    assert int(tot.item()) == want.tokens
    # balanced owners: the collective that picks world-1 splitters from a finished job, with a
    # stand-in for the context (the rows this rank would hold under partition-mod-world ownership)
    def sort_word(key):
        return int.from_bytes(key[:7].ljust(7, b"\0"), "big") << 8 | min(len(key), 8)

    my_rows = [(int(p), k) for p, k in zip(want.part, want.keys()) if parallel.owner_of_partition(int(p), world) == rank]

    class FakeContext:
        def __init__(self):
            self.world_size, self.rank, self.rows, self.installed = world, rank, my_rows, None

        def job_sync(self):
            return None

        def rows_part_begin(self):
            return np.searchsorted(np.array([p for p, _ in self.rows], dtype=np.int64), np.arange(R + 1)).astype(np.uint64)

        def row_sort_word(self, row):
            p, k = self.rows[row]
            return p, sort_word(k)

        def set_owners(self, splitters):
            self.installed = list(splitters)

    fake = FakeContext()
    splitters = parallel.balance_owners(fake)
    assert fake.installed == splitters and len(splitters) == world - 1 and splitters == sorted(splitters)
    everyone = [None] * world
    dist.all_gather_object(everyone, splitters)
    assert all(e == splitters for e in everyone)                      # every rank installs the same cut points
    owner = [sum(1 for sp in splitters if sp <= (int(p), sort_word(k))) for p, k in zip(want.part, want.keys())]
    assert owner == sorted(owner)                                     # contiguous pieces of the output order
    share = np.bincount(owner, minlength=world)
    assert share.max() - share.min() <= 16, share                     # equal shares (keys equal in 7 bytes stay together)
    dist.barrier()
    dist.destroy_process_group()
    print(f"rank {rank}: ok, {len(merged)} keys owned")


if __name__ == "__main__":
    main()
