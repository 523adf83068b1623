"""Does a device->host copy running beside the host->device copies of the splits slow them down?"""
import time, numpy as np, torch
dev = torch.device("cuda", 0)
n, sz = 197, 1_495_000
src = [torch.empty(sz, dtype=torch.uint8).pin_memory() for _ in range(n)]
dst = torch.empty(n * (sz + 4096), dtype=torch.uint8, device=dev)
back_d = torch.empty(16_000_000, dtype=torch.uint8, device=dev)
back_h = torch.empty(16_000_000, dtype=torch.uint8).pin_memory()
sa, sb = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
def run(duplex, reps):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(reps):
        if duplex:
            with torch.cuda.stream(sb): back_h.copy_(back_d, non_blocking=True)
        off = 0
        with torch.cuda.stream(sa):
            for t in src:
                dst[off:off + sz].copy_(t, non_blocking=True); off += sz + 4096
        if not duplex:
            with torch.cuda.stream(sa): back_h.copy_(back_d, non_blocking=True)
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3
for d in (False, True, False, True):
    run(d, 2); print("D2H beside the H2D copies" if d else "D2H after the H2D copies ", f"{run(d, 5):.3f} ms per round")
