"""The owner's side of an N-rank job on ONE GPU, for profiling: N contexts on device 0, every rank maps its
own config-2 shard (197 splits), records cross through the fixed-slot transport with the all-to-all done
as device copies, and rank 0's reduce side -- the kernels an owner runs at N GPUs, on the data an owner
gets there (every key in up to N copies) -- is timed with CUDA events.  Run it under ncu for a launch
list of those kernels:

    python scripts/owner_side.py --ranks 8 [--jobs 6]
    ncu --metrics gpu__time_duration.sum --clock-control none -k regex:msd_ ... python scripts/owner_side.py --ranks 8
"""
import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from lua_mapreduce_1_b200 import _native, corpus, parallel  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ranks", type=int, default=8)
    ap.add_argument("--jobs", type=int, default=6)
    a = ap.parse_args()
    W, R = a.ranks, 10
    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    lines = corpus.split_lines(1)
    ctxs = []
    for r in range(W):
        splits, _ = corpus.generate(lines=lines, first_split=r * len(lines))
        total = sum(len(s) for s in splits)
        c = _native.Context(R, total + 64 * len(splits) + (1 << 16), max_keys=1 << 21, table_slots=1 << 21, device=0,
                            stream=stream.cuda_stream, world_size=W, rank=r)
        for i, s in enumerate(splits):
            c.map_split(i, s)
        ctxs.append(c)
    slot = 24 << 20

    def u8(ptr, n):
        return parallel.device_bytes(ptr, n, dev)

    def job(first):
        for c in ctxs:
            if not first:
                c.job_begin(keep_input=True)
            c.map_finish()
        send = [u8(c.shuffle_pack_fixed(slot), W * slot) for c in ctxs]
        recv = [u8(c.shuffle_recv_buffer(W * slot), W * slot) for c in ctxs]
        for s in range(W):
            for d in range(W):
                recv[d][s * slot:(s + 1) * slot].copy_(send[s][d * slot:(d + 1) * slot])
        for c in ctxs:
            c.shuffle_import_fixed(slot)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ctxs[0].reduce_all()
        e1.record()
        for c in ctxs[1:]:
            c.reduce_all()
        return e0, e1

    # job 1 with owner = partition mod ranks, then equal shares of the output order (as bench.py does)
    job(True)
    rows = [c.rows() for c in ctxs]
    order = []
    for r, rw in enumerate(rows):
        keys = rw.keys()
        order += [(int(rw.part[i]), keys[i]) for i in range(rw.n)]
    order.sort()
    from helpers import sort_word
    cuts = [order[len(order) * j // W] for j in range(1, W)]
    splitters = [(p, sort_word(k)) for p, k in cuts]
    for c in ctxs:
        c.set_owners(splitters)
    times = []
    for j in range(a.jobs):
        e0, e1 = job(False)
        info = ctxs[0].job_sync()
        for c in ctxs[1:]:
            c.job_sync()
        times.append(e0.elapsed_time(e1) * 1e3)
    print(f"ranks {W}: owner 0 reduces {int(info.n_keys)} rows; reduce side per job (us): " + " ".join(f"{t:.0f}" for t in times)
          + f"; flags {int(info.flags):#x}")
    L = _native.load()
    if hasattr(L, "mr_debug_stamps"):        # a -DMRGPU_UNIT_TIMING=<unit> build: that unit's phases on owner 0
        import ctypes as C
        L.mr_debug_stamps.restype = C.POINTER(C.c_longlong)
        L.mr_debug_stamps.argtypes = [C.c_void_p]
        d = L.mr_debug_stamps(ctxs[0]._h)
        st = [d[i] for i in range(16)]
        names = ["start", "ticket+range", "load", "sort", "gather", "ties", "heads+merge+stage", "look-back", "rows", "copy-out"]
        print("unit phases (us): starts %.2f after unit 0; " % ((st[0] - st[15]) / 1e3) +
              "  ".join(f"{n} +{(st[k] - st[k - 1]) / 1e3:.2f}" for k, n in enumerate(names) if k) + f" | total {(st[9] - st[0]) / 1e3:.2f}")
    for c in ctxs:
        c.close()


if __name__ == "__main__":
    main()
