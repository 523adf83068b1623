"""Variant libraries for scripts/owner_side.py: counting-sort thresholds and unit phase stamps."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "build", "variants")
V = {"cs48": ["-DMRGPU_COUNT_SORT_MAX=48"], "cs192": ["-DMRGPU_COUNT_SORT_MAX=192"], "cs1024": ["-DMRGPU_COUNT_SORT_MAX=1024"],
     "ut5": ["-DMRGPU_UNIT_TIMING=5"], "ut60": ["-DMRGPU_UNIT_TIMING=60"], "ut400": ["-DMRGPU_UNIT_TIMING=400"], "ut900": ["-DMRGPU_UNIT_TIMING=900"]}
if sys.argv[1] == "build":
    from lua_mapreduce_1_b200 import build as b
    os.makedirs(OUT, exist_ok=True)
    srcs, _ = b.gpu_sources()
    for n, f in V.items():
        subprocess.check_call([b._nvcc()] + b.NVCC_FLAGS + f + ["-o", os.path.join(OUT, f"libmrgpu_{n}.so")] + srcs)
else:
    for n in V:
        print("==", n, flush=True)
        subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "owner_side.py"), "--ranks", sys.argv[2], "--jobs", "4"],
                       env=dict(os.environ, MRGPU_LIB=os.path.join(OUT, f"libmrgpu_{n}.so")))
