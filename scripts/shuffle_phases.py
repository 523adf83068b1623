"""Under torchrun: CUDA-event time of each phase of the multi-GPU job (map, pack, all-to-all, import+reduce)."""
import os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lua_mapreduce_1_b200 import _native, corpus, parallel

lr = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
rank, W = dist.get_rank(), dist.get_world_size()
stream = torch.cuda.Stream(device=dev); torch.cuda.set_stream(stream)
splits, tok = corpus.generate(first_split=rank * 197, threads=max(1, os.cpu_count() // W))
total = sum(len(s) for s in splits)
ctx = _native.Context(10, total + 64 * len(splits) + (1 << 16), max_keys=1 << 21, table_slots=1 << 21, device=lr,
                      stream=stream.cuda_stream, world_size=W, rank=rank)
for i, s in enumerate(splits): ctx.map_split(i, s)
ctx.map_finish(); big = parallel.shuffle(ctx); ctx.reduce_all(); ctx.job_sync()
slot = parallel.slot_for(big)
names = ["clear+map", "count+pack", "all_to_all", "import", "reduce"]
acc = np.zeros(len(names)); n = 10
for it in range(n + 3):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(len(names) + 1)]
    dist.barrier(); torch.cuda.synchronize()
    ev[0].record()
    ctx.job_begin(keep_input=True); ctx.map_finish(); ev[1].record()
    tot = W * slot
    send = parallel.device_bytes(ctx.shuffle_pack_fixed(slot), tot, dev); ev[2].record()
    recv = parallel.device_bytes(ctx.shuffle_recv_buffer(tot), tot, dev)
    dist.all_to_all_single(recv, send); ev[3].record()
    ctx.shuffle_import_fixed(slot); ev[4].record()
    ctx.reduce_all(); ev[5].record()
    torch.cuda.synchronize()
    if it >= 3: acc += [ev[i].elapsed_time(ev[i + 1]) for i in range(len(names))]
info = ctx.job_sync()
t = torch.tensor(acc / n, device=dev); mx = t.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
rows = torch.tensor([info.n_keys], device=dev); allrows = [torch.zeros_like(rows) for _ in range(W)]; dist.all_gather(allrows, rows)
if rank == 0:
    print(f"W={W} slot={slot} rows/rank={[int(r.item()) for r in allrows]}")
    for nm, a, b in zip(names, t.tolist(), mx.tolist()): print(f"  {nm:12s} rank0 {a*1e3:8.1f} us   max {b*1e3:8.1f} us")
    print(f"  sum rank0 {t.sum().item()*1e3:.1f} us")
parallel.p2p_connect(ctx, parallel.p2p_slot_for(big))
names = ["clear+map", "shuffle_p2p", "reduce"]
acc = np.zeros(len(names))
for it in range(n + 3):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(len(names) + 1)]
    dist.barrier(); torch.cuda.synchronize()
    ev[0].record()
    ctx.job_begin(keep_input=True); ctx.map_finish(); ev[1].record()
    ctx.shuffle_p2p(); ev[2].record()
    ctx.reduce_all(); ev[3].record()
    torch.cuda.synchronize()
    if it >= 3: acc += [ev[i].elapsed_time(ev[i + 1]) for i in range(len(names))]
t = torch.tensor(acc / n, device=dev); mx = t.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
if rank == 0:
    print(f"peer-to-peer, slot={ctx_slot if 'ctx_slot' in globals() else parallel.p2p_slot_for(big)}")
    for nm, a, b in zip(names, t.tolist(), mx.tolist()): print(f"  {nm:12s} rank0 {a*1e3:8.1f} us   max {b*1e3:8.1f} us")
    print(f"  sum rank0 {t.sum().item()*1e3:.1f} us")
dist.barrier(); ctx.close(); dist.barrier(); dist.destroy_process_group()
