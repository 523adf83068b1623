"""In-tree build of the native libraries (no JIT cache: the .so files travel with the repo).

  libmrgpu.so     CUDA kernels + C ABI (include/mrgpu.h), nvcc for sm_100a only
  libmrcorpus.so  synthetic corpus generator (host C)
  examples/c_host/wordcount_host   a worker written in C against include/mrgpu.h alone
"""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB_GPU = os.path.join(HERE, "libmrgpu.so")
LIB_CORPUS = os.path.join(HERE, "libmrcorpus.so")
C_HOST = os.path.join(HERE, "examples", "c_host", "wordcount_host")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-shared", "-cudart", "shared",
    "-I", os.path.join(ROOT, "include"), "-I", CSRC,
]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _nvcc():
    for c in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found: the CUDA library cannot be built")


def gpu_sources():
    srcs = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith(".cu")]
    deps = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith((".cuh", ".h"))]
    deps.append(os.path.join(ROOT, "include", "mrgpu.h"))
    return srcs, deps


def build_gpu(force=False, verbose=False, extra=()):
    srcs, deps = gpu_sources()
    if not force and not _newer(LIB_GPU, srcs + deps):
        return LIB_GPU
    cmd = [_nvcc()] + NVCC_FLAGS + list(extra) + ["-o", LIB_GPU] + srcs
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    subprocess.check_call(cmd)
    return LIB_GPU


def build_corpus(force=False):
    src = os.path.join(CSRC, "corpus.c")
    if not force and not _newer(LIB_CORPUS, [src]):
        return LIB_CORPUS
    subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", "-std=gnu11", "-o", LIB_CORPUS, src])
    return LIB_CORPUS


def build_c_host(force=False):
    """The C worker links libmrgpu.so by relative rpath, so the pair travels together."""
    src = C_HOST + ".c"
    if not force and not _newer(C_HOST, [src, LIB_GPU, os.path.join(ROOT, "include", "mrgpu.h")]):
        return C_HOST
    subprocess.check_call(["gcc", "-O2", "-std=gnu11", "-Wall", "-I", os.path.join(ROOT, "include"), "-o", C_HOST, src,
                           "-L", HERE, "-lmrgpu", "-Wl,-rpath,$ORIGIN/../.."])
    return C_HOST


def build_all(force=False, verbose=False):
    return build_gpu(force, verbose), build_corpus(force), build_c_host(force)


if __name__ == "__main__":
    import sys
    print(build_all(force="--force" in sys.argv, verbose="-v" in sys.argv))
