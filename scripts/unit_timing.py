"""Phase stamps (globaltimer, ns) of single units of msd_unit_kernel on config 2: libraries built with
-DMRGPU_UNIT_TIMING=<unit> (`build` here, `run` on the GPU box)."""
import ctypes as C, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "build", "variants")
UNITS = [5, 300, 590, 700, 1000]
NAMES = ["start", "ticket+range", "load", "sort", "gather", "ties", "heads+merge+stage", "look-back", "rows", "copy-out"]

def build():
    from lua_mapreduce_1_b200 import build as b
    os.makedirs(OUT, exist_ok=True)
    srcs, _ = b.gpu_sources()
    for u in UNITS:
        subprocess.check_call([b._nvcc()] + b.NVCC_FLAGS + [f"-DMRGPU_UNIT_TIMING={u}", "-o", os.path.join(OUT, f"libmrgpu_ut{u}.so")] + srcs)

def one(u):
    from lua_mapreduce_1_b200 import _native, corpus
    splits, _ = corpus.generate()
    ctx = _native.Context(10, sum(len(s) for s in splits) + (1 << 20), max_keys=1 << 20, table_slots=1 << 21)
    for i, s in enumerate(splits): ctx.map_split(i, s)
    ctx.job_sync()
    for _ in range(3):
        ctx.job_begin(keep_input=True); ctx.map_finish(); ctx.reduce_all(); ctx.job_sync()
    L = _native.load(); L.mr_debug_stamps.restype = C.POINTER(C.c_longlong); L.mr_debug_stamps.argtypes = [C.c_void_p]
    d = L.mr_debug_stamps(ctx._h); st = [d[i] for i in range(16)]
    t0 = st[15]
    print(f"unit {u}: starts {((st[0] - t0) / 1e3):7.2f} us after unit 0;", "  ".join(f"{n} +{(st[k] - st[k - 1]) / 1e3:.2f}" for k, n in enumerate(NAMES) if k), f"| total {(st[9] - st[0]) / 1e3:.2f} us, ends at {(st[9] - t0) / 1e3:.2f}")

def run():
    for u in UNITS:
        subprocess.run([sys.executable, os.path.abspath(__file__), "one", str(u)], env=dict(os.environ, MRGPU_LIB=os.path.join(OUT, f"libmrgpu_ut{u}.so")))

if __name__ == "__main__":
    {"build": build, "run": run, "one": lambda: one(int(sys.argv[2]))}[sys.argv[1]]()
