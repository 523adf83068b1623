"""Where the end-to-end step goes: raw pinned H2D rate, the staged job, and its tail (GPU box only)."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lua_mapreduce_1_b200 import _native, corpus

dev = torch.device("cuda", 0)
splits, tok = corpus.generate(threads=os.cpu_count())
total = sum(len(s) for s in splits)
pinned = [torch.from_numpy(s).pin_memory() for s in splits]
big = torch.empty(total, dtype=torch.uint8).pin_memory()
dst = torch.empty(total + 4096 * len(splits), dtype=torch.uint8, device=dev)
s = torch.cuda.Stream(device=dev)
def timed(fn, n=5):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3
with torch.cuda.stream(s):
    ms = timed(lambda: dst[:total].copy_(big, non_blocking=True))
    print(f"one {total/1e6:.1f} MB pinned copy: {ms:.3f} ms = {total/ms/1e6:.1f} GB/s")
    def many():
        off = 0
        for t in pinned:
            dst[off:off + t.numel()].copy_(t, non_blocking=True); off += (t.numel() + 4095) // 4096 * 4096
    ms = timed(many)
    print(f"{len(pinned)} pinned copies: {ms:.3f} ms = {total/ms/1e6:.1f} GB/s")
stream = torch.cuda.Stream(device=dev); torch.cuda.set_stream(stream)
ctx = _native.Context(10, total + 64 * len(splits) + (1 << 16), max_keys=1 << 20, table_slots=1 << 21, stream=stream.cuda_stream)
def job(sync_points=False):
    ctx.job_begin()
    t0 = time.perf_counter()
    for i, t in enumerate(pinned): ctx.map_split_ptr(i, t.data_ptr(), t.numel())
    t1 = time.perf_counter()
    if sync_points: torch.cuda.synchronize()
    t2 = time.perf_counter()
    ctx.map_finish()
    if sync_points: torch.cuda.synchronize()
    t3 = time.perf_counter()
    ctx.reduce_all()
    if sync_points: torch.cuda.synchronize()
    t4 = time.perf_counter()
    rows = ctx.rows(OUT[0])
    t5 = time.perf_counter()
    return [(b - a) * 1e3 for a, b in ((t0, t1), (t1, t2), (t2, t3), (t3, t4), (t4, t5))]
OUT = [None]
job(); r0 = ctx.rows()
def pb(n, dt): return torch.empty(max(int(n), 1), dtype=dt).pin_memory().numpy()[:int(n)]
OUT[0] = _native.Rows(pb(r0.n, torch.int32).view(np.uint32), pb(r0.n, torch.int64).view(np.uint64), pb(r0.n + 1, torch.int64).view(np.uint64),
                      pb(len(r0.key_bytes), torch.uint8), pb(11, torch.int64).view(np.uint64))
job()
for sp in (False, True):
    acc = np.zeros(5)
    n = 5
    w0 = time.perf_counter()
    for _ in range(n): acc += job(sp)
    w = (time.perf_counter() - w0) / n * 1e3
    print(f"sync_points={sp}: step {w:.3f} ms; stage calls {acc[0]/n:.3f}, wait copies {acc[1]/n:.3f}, map_finish {acc[2]/n:.3f}, reduce {acc[3]/n:.3f}, rows {acc[4]/n:.3f}")
for mb in (4, 8, 16, 32, 64):
    os.environ["MRGPU_AUTO_MAP_MB"] = str(mb)
    c2 = _native.Context(10, total + 64 * len(splits) + (1 << 16), max_keys=1 << 20, table_slots=1 << 21, stream=stream.cuda_stream)
    def j2():
        c2.job_begin()
        for i, t in enumerate(pinned): c2.map_split_ptr(i, t.data_ptr(), t.numel())
        c2.map_finish(); c2.reduce_all(); c2.rows(OUT[0])
    j2(); j2()
    t0 = time.perf_counter()
    for _ in range(5): j2()
    print(f"auto map every {mb} MB: {(time.perf_counter()-t0)/5*1e3:.3f} ms")
    c2.close()
