"""Per-phase cycles of one warp of the map kernel over a whole resident launch (library built with
-DMRGPU_MAP_TIMING)."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from lua_mapreduce_1_b200 import _native, corpus
splits, _ = corpus.generate()
ctx = _native.Context(10, sum(len(s) for s in splits) + (1 << 20), max_keys=1 << 20, table_slots=1 << 21)
for i, s in enumerate(splits): ctx.map_split(i, s)
ctx.job_sync()
for _ in range(2):
    ctx.job_begin(keep_input=True); ctx.map_finish(); ctx.reduce_all(); ctx.job_sync()
L = _native.load(); L.mr_debug_stamps.restype = C.POINTER(C.c_longlong); L.mr_debug_stamps.argtypes = [C.c_void_p]
d = L.mr_debug_stamps(ctx._h); st = [d[i] for i in range(8)]
names = ["1 loads+stage+mask", "2 starts/classes/scan/lists", "3a short: key+CTA table", "3b medium: key+queue",
         "drains (global table)", "final drain", "-", "loop overhead"]
tot = sum(st)
for nme, v in zip(names, st): print(f"{nme:30s} {v:10d} cycles {v / tot * 100:5.1f}%")
print("total", tot, "cycles =", tot / 1.965e3, "us")
x = [d[i] for i in range(8, 12)]
print(f"table init {x[0]} cycles; wait for the CTA's other warps {x[1]}; flush (this warp) {x[2]}; keys in the CTA table {x[3]}")
