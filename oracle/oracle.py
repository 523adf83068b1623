"""ctypes wrapper around oracle/libmr_oracle.so (TEST INFRASTRUCTURE ONLY).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs import this module; the product package never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libmr_oracle.so")


def build(force=False):
    src = os.path.join(_HERE, "mr_oracle.c")
    hdr = os.path.join(_HERE, "mr_oracle.h")
    if (not force and os.path.exists(_SO)
            and os.path.getmtime(_SO) >= max(os.path.getmtime(src), os.path.getmtime(hdr))):
        return _SO
    subprocess.check_call(["make", "-C", _HERE, "-B", "libmr_oracle.so"],
                          stdout=subprocess.DEVNULL)
    return _SO


class _Result(C.Structure):
    _fields_ = [("n", C.c_uint64), ("tokens", C.c_uint64),
                ("part", C.POINTER(C.c_uint32)), ("count", C.POINTER(C.c_uint64)),
                ("key_off", C.POINTER(C.c_uint64)), ("key_bytes", C.POINTER(C.c_uint8)),
                ("key_bytes_len", C.c_uint64)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        L.mro_fnv_lua52.restype = C.c_uint32
        L.mro_fnv_lua52.argtypes = [C.c_char_p, C.c_uint64]
        L.mro_fnv_lua52_int.restype = C.c_uint32
        L.mro_fnv_lua52_int.argtypes = [C.c_char_p, C.c_uint64]
        L.mro_partition.restype = C.c_uint32
        L.mro_partition.argtypes = [C.c_char_p, C.c_uint64, C.c_uint32]
        L.mro_key_cmp.restype = C.c_int
        L.mro_key_cmp.argtypes = [C.c_char_p, C.c_uint64, C.c_char_p, C.c_uint64]
        L.mro_is_space.restype = C.c_int
        L.mro_is_space.argtypes = [C.c_ubyte]
        for f in (L.mro_run_naive, L.mro_run_staged, L.mro_run_bigram):
            f.restype = C.POINTER(_Result)
        L.mro_run_bigram.argtypes = [C.POINTER(C.c_void_p), C.POINTER(C.c_uint64), C.c_uint32, C.c_uint32]
        L.mro_run_naive.argtypes = [C.POINTER(C.c_void_p), C.POINTER(C.c_uint64), C.c_uint32,
                                    C.c_uint32]
        L.mro_run_staged.argtypes = [C.POINTER(C.c_void_p), C.POINTER(C.c_uint64), C.c_uint32,
                                     C.c_uint32, C.c_int]
        L.mro_groupby_u32.restype = C.POINTER(_Result)
        L.mro_groupby_u32.argtypes = [C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_uint64),
                                      C.c_uint32, C.c_uint32, C.c_int]
        L.mro_result_free.argtypes = [C.POINTER(_Result)]
        L.mro_heap_sort_u64.argtypes = [C.POINTER(C.c_uint64), C.c_uint32]
        L.mro_merge_kat.restype = C.c_uint32
        L.mro_merge_kat.argtypes = [C.POINTER(C.POINTER(C.c_uint64)),
                                    C.POINTER(C.POINTER(C.c_uint64)),
                                    C.POINTER(C.c_uint32), C.c_uint32,
                                    C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.mro_max_threads.restype = C.c_int
        _lib = L
    return _lib


class Rows:
    """(partition, key, count) rows in the reference's final order."""

    def __init__(self, part, count, key_off, key_bytes, tokens=0):
        self.part = np.asarray(part, dtype=np.uint32)
        self.count = np.asarray(count, dtype=np.uint64)
        self.key_off = np.asarray(key_off, dtype=np.uint64)
        self.key_bytes = np.asarray(key_bytes, dtype=np.uint8)
        self.tokens = int(tokens)

    @property
    def n(self):
        return len(self.part)

    def key(self, i):
        return self.key_bytes[int(self.key_off[i]):int(self.key_off[i + 1])].tobytes()

    def keys(self):
        kb = self.key_bytes.tobytes()
        off = self.key_off
        return [kb[int(off[i]):int(off[i + 1])] for i in range(self.n)]

    def as_dict(self):
        return dict(zip(self.keys(), (int(c) for c in self.count)))

    def same_as(self, other):
        return (self.n == other.n and np.array_equal(self.part, other.part)
                and np.array_equal(self.count, other.count)
                and np.array_equal(self.key_off, other.key_off)
                and np.array_equal(self.key_bytes, other.key_bytes))


def _splits_args(splits):
    bufs = [np.frombuffer(s, dtype=np.uint8) if not isinstance(s, np.ndarray) else s
            for s in splits]
    n = len(bufs)
    ptrs = (C.c_void_p * max(n, 1))(*[b.ctypes.data if b.size else None for b in bufs])
    lens = (C.c_uint64 * max(n, 1))(*[b.size for b in bufs])
    return bufs, ptrs, lens, n


def _take(res):
    r = res.contents
    n = int(r.n)
    out = Rows(np.ctypeslib.as_array(r.part, (n,)).copy() if n else [],
               np.ctypeslib.as_array(r.count, (n,)).copy() if n else [],
               np.ctypeslib.as_array(r.key_off, (n + 1,)).copy(),
               np.ctypeslib.as_array(r.key_bytes, (int(r.key_bytes_len),)).copy()
               if r.key_bytes_len else [],
               r.tokens)
    lib().mro_result_free(res)
    return out


def run_naive(splits, num_reducers):
    """misc/naive.lua over the concatenation of splits (tokens never span splits)."""
    bufs, ptrs, lens, n = _splits_args(splits)
    return _take(lib().mro_run_naive(ptrs, lens, n, num_reducers))


def run_staged(splits, num_reducers, nthreads=1):
    """job.lua map jobs -> sorted runs -> utils.merge_iterator reduce jobs."""
    bufs, ptrs, lens, n = _splits_args(splits)
    return _take(lib().mro_run_staged(ptrs, lens, n, num_reducers, nthreads))


def run_bigram(splits, num_reducers):
    """Bigram count (BASELINE config 4): keys `w_i .. " " .. w_{i+1}` inside a line; .tokens = pairs emitted."""
    bufs, ptrs, lens, n = _splits_args(splits)
    return _take(lib().mro_run_bigram(ptrs, lens, n, num_reducers))


def groupby_u32(splits, num_reducers, op=0):
    """splits: list of (keys uint32 array, values uint32 array).  Group-by-sum on number keys
    (BASELINE config 5); key bytes in the rows are 4 big-endian bytes."""
    ks = [np.ascontiguousarray(k, dtype=np.uint32) for k, _ in splits]
    vs = [np.ascontiguousarray(v, dtype=np.uint32) for _, v in splits]
    n = len(ks)
    kp = (C.c_void_p * max(n, 1))(*[a.ctypes.data for a in ks])
    vp = (C.c_void_p * max(n, 1))(*[a.ctypes.data for a in vs])
    lens = (C.c_uint64 * max(n, 1))(*[a.size for a in ks])
    return _take(lib().mro_groupby_u32(kp, vp, lens, n, num_reducers, op))


def fnv_lua52(key: bytes) -> int:
    return lib().mro_fnv_lua52(key, len(key))


def fnv_lua52_int(key: bytes) -> int:
    return lib().mro_fnv_lua52_int(key, len(key))


def partition(key: bytes, num_reducers: int) -> int:
    return lib().mro_partition(key, len(key), num_reducers)


def key_cmp(a: bytes, b: bytes) -> int:
    return lib().mro_key_cmp(a, len(a), b, len(b))


def heap_sort(values):
    arr = (C.c_uint64 * len(values))(*values)
    lib().mro_heap_sort_u64(arr, len(values))
    return list(arr)


def merge_kat(runs):
    """runs: list of lists of (numeric key, number of values).  utils.lua:360-380."""
    n = len(runs)
    karrs = [(C.c_uint64 * max(len(r), 1))(*[k for k, _ in r]) for r in runs]
    carrs = [(C.c_uint64 * max(len(r), 1))(*[c for _, c in r]) for r in runs]
    kp = (C.POINTER(C.c_uint64) * n)(*[C.cast(a, C.POINTER(C.c_uint64)) for a in karrs])
    cp = (C.POINTER(C.c_uint64) * n)(*[C.cast(a, C.POINTER(C.c_uint64)) for a in carrs])
    lens = (C.c_uint32 * n)(*[len(r) for r in runs])
    total = sum(len(r) for r in runs)
    ok = (C.c_uint64 * max(total, 1))()
    oc = (C.c_uint64 * max(total, 1))()
    m = lib().mro_merge_kat(kp, cp, lens, n, ok, oc)
    return [(int(ok[i]), int(oc[i])) for i in range(m)]


def max_threads():
    return lib().mro_max_threads()
