"""H2D issue-vs-completion probe: is cudaMemcpyAsync from pinned memory returning early? do two streams help?"""
import os, sys, time
import numpy as np, torch
dev = torch.device("cuda", 0)
n, sz = 197, 1_495_000
src = [torch.empty(sz, dtype=torch.uint8).pin_memory() for _ in range(n)]
dst = torch.empty(n * (sz + 4096), dtype=torch.uint8, device=dev)
streams = [torch.cuda.Stream(device=dev) for _ in range(4)]
def run(k, chunks=1):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    off = 0
    for i, t in enumerate(src):
        step = (sz + chunks - 1) // chunks
        for c in range(chunks):
            a, b = c * step, min(sz, (c + 1) * step)
            with torch.cuda.stream(streams[(i * chunks + c) % k]):
                dst[off + a:off + b].copy_(t[a:b], non_blocking=True)
        off += sz + 4096
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    return (t1 - t0) * 1e3, (t2 - t0) * 1e3
for k in (1, 2, 3, 4):
    run(k)
    r = [run(k) for _ in range(5)]
    print(f"{k} stream(s): issue {np.mean([a for a, _ in r]):.3f} ms, done {np.mean([b for _, b in r]):.3f} ms = {n*sz/np.mean([b for _, b in r])/1e6:.1f} GB/s")
for ch in (2, 4):
    run(2, ch)
    r = [run(2, ch) for _ in range(5)]
    print(f"2 streams, {ch} chunks per split: issue {np.mean([a for a, _ in r]):.3f} ms, done {np.mean([b for _, b in r]):.3f} ms")
