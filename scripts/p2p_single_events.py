"""World of one, config-2 shard: CUDA-event time of shuffle_p2p alone (no peer to wait for)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lua_mapreduce_1_b200 import _native, corpus
splits, tok = corpus.generate(threads=os.cpu_count())
total = sum(len(s) for s in splits)
st = torch.cuda.Stream(); torch.cuda.set_stream(st)
ctx = _native.Context(10, total + 64 * len(splits) + (1 << 16), max_keys=1 << 21, table_slots=1 << 21, stream=st.cuda_stream)
for i, s in enumerate(splits): ctx.map_split(i, s)
ctx.map_finish()
ctx.p2p_attach([ctx.p2p_setup(24 << 20)])
acc = np.zeros(3); n = 10
for it in range(n + 3):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    torch.cuda.synchronize()
    ev[0].record()
    if it: ctx.job_begin(keep_input=True); ctx.map_finish()
    ev[1].record(); ctx.shuffle_p2p(); ev[2].record(); ctx.reduce_all(); ev[3].record()
    torch.cuda.synchronize()
    if it >= 3: acc += [ev[i].elapsed_time(ev[i + 1]) for i in range(3)]
print("map %.1f us  shuffle_p2p %.1f us  reduce %.1f us" % tuple(acc / n * 1e3))
ctx.close()
