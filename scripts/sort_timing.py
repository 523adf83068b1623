"""Phase timing of one radix-pass tile (needs a library built with -DMRGPU_SORT_TIMING)."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from lua_mapreduce_1_b200 import _native, corpus
splits, _ = corpus.generate()
ctx = _native.Context(10, sum(len(s) for s in splits) + (1 << 20), max_keys=1 << 20, table_slots=1 << 21)
for i, s in enumerate(splits): ctx.map_split(i, s)
ctx.job_sync()
for _ in range(3):
    ctx.job_begin(keep_input=True); ctx.map_finish(); ctx.reduce_all(); ctx.job_sync()
L = _native.load(); L.mr_debug_stamps.restype = C.POINTER(C.c_longlong); L.mr_debug_stamps.argtypes = [C.c_void_p]
d = L.mr_debug_stamps(ctx._h); st = [d[i] for i in range(9)]
names = ["loads issued+setup+sync", "digits (wait loads)", "ballot+rank loop", "barrier", "bins+publish", "look-back", "scans+syncs", "reorder+scatter"]
for i, nme in enumerate(names): print(f"{nme:22s} {st[i+1]-st[i]:8d} cycles")
print("total", st[8] - st[0])
