"""World of one, config-2 shard, peer-to-peer shuffle against the context's own window: for a per-kernel launch list."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lua_mapreduce_1_b200 import _native, corpus
splits, tok = corpus.generate(threads=os.cpu_count())
total = sum(len(s) for s in splits)
ctx = _native.Context(10, total + 64 * len(splits) + (1 << 16), max_keys=1 << 21, table_slots=1 << 21)
for i, s in enumerate(splits): ctx.map_split(i, s)
ctx.map_finish()
ctx.p2p_attach([ctx.p2p_setup(24 << 20)])
for it in range(3):
    if it: ctx.job_begin(keep_input=True); ctx.map_finish()
    ctx.shuffle_p2p(); ctx.reduce_all(); info = ctx.job_sync()
print(info.n_keys, info.n_tokens, tok)
ctx.close()
