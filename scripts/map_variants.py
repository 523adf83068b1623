"""Builds variant libraries of the map kernel (-DMRGPU_V_* switches of csrc/map_kernel.cuh) and, on a
GPU, times each on config 2 (resident jobs, CUDA events of mr_stats) and checks its rows against the
default build's.  `python scripts/map_variants.py build` here, `... run` on the GPU box."""
import hashlib
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "build", "variants")
VARIANTS = {
    "base": [],
    "plainlds": ["-DMRGPU_V_PLAINLDS"],
    "onebyte": ["-DMRGPU_V_ONEBYTE"],
    "queuemed": ["-DMRGPU_V_QUEUEMED"],
    "plainlds_onebyte": ["-DMRGPU_V_PLAINLDS", "-DMRGPU_V_ONEBYTE"],
    "asyncclaim": ["-DMRGPU_V_ASYNCCLAIM"],
    # warps per SM (registers per thread follow from the launch bound) x CTA-table size
        "w32_12": ["-DMRGPU_V_WARPS=32", "-DMRGPU_V_LOG2=12"],
    "w32_11": ["-DMRGPU_V_WARPS=32", "-DMRGPU_V_LOG2=11"],
    "w32_10": ["-DMRGPU_V_WARPS=32", "-DMRGPU_V_LOG2=10"],
    "w24_12": ["-DMRGPU_V_WARPS=24", "-DMRGPU_V_LOG2=12"],
    "w16_12x2": ["-DMRGPU_V_WARPS=16", "-DMRGPU_V_LOG2=12", "-DMRGPU_V_CTAS=2"],
    "w16_11x2": ["-DMRGPU_V_WARPS=16", "-DMRGPU_V_LOG2=11", "-DMRGPU_V_CTAS=2"],
    "w20_12x2": ["-DMRGPU_V_WARPS=20", "-DMRGPU_V_LOG2=12", "-DMRGPU_V_CTAS=2"],
    "w20_11x2": ["-DMRGPU_V_WARPS=20", "-DMRGPU_V_LOG2=11", "-DMRGPU_V_CTAS=2"],
    "w24_11x2": ["-DMRGPU_V_WARPS=24", "-DMRGPU_V_LOG2=11", "-DMRGPU_V_CTAS=2"],
    "w24_10x2": ["-DMRGPU_V_WARPS=24", "-DMRGPU_V_LOG2=10", "-DMRGPU_V_CTAS=2"],
}
if os.environ.get("MAP_VARIANTS"):
    VARIANTS = {k: VARIANTS[k] for k in os.environ["MAP_VARIANTS"].split(",")}


def build():
    from lua_mapreduce_1_b200 import build as b
    os.makedirs(OUT, exist_ok=True)
    srcs, _ = b.gpu_sources()
    for name, flags in VARIANTS.items():
        want = [("32", "13")]
        if any("V_WARPS" in f for f in flags):
            want = [(flags[0].split("=")[1], flags[1].split("=")[1])]
        lib = os.path.join(OUT, f"libmrgpu_{name}.so")
        cmd = [b._nvcc()] + b.NVCC_FLAGS + flags + ["-Xptxas=-v", "-o", lib] + srcs
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode:
            print(r.stderr[-3000:]); raise SystemExit(1)
        lines = r.stderr.splitlines()
        for i, l in enumerate(lines):
            if "map_combine_kernel" in l and "Compiling" in l and any(f"ILi{w}ELi{g}E" in l for w, g in want):
                print(name, "|", lines[i + 2].strip(), "|", lines[i + 1].strip())


def one(name):
    import numpy as np
    from lua_mapreduce_1_b200 import _native, corpus
    splits, tok = corpus.generate()
    slots = int(os.environ.get("MAP_TABLE_SLOTS", 1 << 21))      # the global table: 2^21 slots of 32 bytes by default
    ctx = _native.Context(10, sum(len(s) for s in splits) + (1 << 20), max_keys=1 << 20, table_slots=slots)
    for i, s in enumerate(splits):
        ctx.map_split(i, s)
    ctx.map_finish(); ctx.reduce_all()
    info = ctx.job_sync()
    assert info.n_tokens == tok
    rows = ctx.rows()
    h = hashlib.sha256()
    for a in (rows.part, rows.count, rows.key_off, rows.key_bytes):
        h.update(np.ascontiguousarray(a).tobytes())
    ms, rs = [], []
    for _ in range(12):
        ctx.job_begin(keep_input=True); ctx.map_finish(); ctx.reduce_all(); ctx.job_sync()
        ms.append(ctx.stats().map_ms); rs.append(ctx.stats().reduce_ms)
    ms = sorted(ms[2:]); rs = sorted(rs[2:])
    print(json.dumps({"variant": name, "map_ms_median": ms[len(ms) // 2], "map_ms_min": ms[0], "reduce_ms_median": rs[len(rs) // 2],
                      "table_slots": slots, "rows_sha": h.hexdigest()[:16]}))


def run():
    for name in VARIANTS:
        env = dict(os.environ, MRGPU_LIB=os.path.join(OUT, f"libmrgpu_{name}.so"))
        subprocess.run([sys.executable, os.path.abspath(__file__), "one", name], env=env)


if __name__ == "__main__":
    {"build": build, "run": run, "one": lambda: one(sys.argv[2])}[sys.argv[1]]()
