"""The device helpers of csrc/common.cuh, compiled for the host (tests/host_helpers.cu): the
partitioner, the sort words and the owner map that the kernels run, checked on CPU against the
oracle's binary64 partitioner, the library's host-side mr_partition_of, and Python's own byte-string
order (Lua's `<` on strings in the C locale).  No GPU needed: the device build of the same source is
byte-for-byte what it was before these functions became __host__ __device__ (checked by comparing
the SASS when that was done)."""
import bisect
import ctypes as C
import os
import random
import subprocess

import pytest

from conftest import ROOT
from lua_mapreduce_1_b200 import _native
from oracle import oracle as O

SRC = os.path.join(ROOT, "tests", "host_helpers.cu")
OUT = os.path.join(ROOT, "tests", "_build", "libhost_helpers.so")


@pytest.fixture(scope="module")
def lib():
    deps = [SRC, os.path.join(ROOT, "lua-mapreduce-1_b200", "csrc", "common.cuh")]
    if not os.path.exists(OUT) or any(os.path.getmtime(d) > os.path.getmtime(OUT) for d in deps):
        os.makedirs(os.path.dirname(OUT), exist_ok=True)
        nvcc = "/usr/local/cuda/bin/nvcc"
        subprocess.check_call([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-O2", "-std=c++17", "-Xcompiler", "-fPIC",
                               "-shared", "-I", os.path.join(ROOT, "lua-mapreduce-1_b200", "csrc"),
                               "-I", os.path.join(ROOT, "include"), "-o", OUT, SRC])
    L = C.CDLL(OUT)
    L.th_fnv.restype = C.c_uint32
    L.th_fnv.argtypes = [C.c_char_p, C.c_uint64]
    L.th_partition_of.restype = C.c_uint32
    L.th_partition_of.argtypes = [C.c_char_p, C.c_uint64, C.c_uint32]
    L.th_sort_word.restype = C.c_uint64
    L.th_sort_word.argtypes = [C.c_char_p, C.c_uint64, C.c_uint32]
    L.th_owner_of.restype = C.c_uint32
    L.th_owner_of.argtypes = [C.c_uint32, C.POINTER(C.c_uint32), C.POINTER(C.c_uint64), C.c_uint32, C.c_uint32, C.c_uint64]
    return L


def random_keys(n, seed, max_len=40):
    rnd = random.Random(seed)
    alphabets = [bytes(range(256)), bytes(range(0x61, 0x7b)), b"ab", bytes(range(0x80, 0x100)), b"\x00\xff"]
    out = []
    for _ in range(n):
        a = rnd.choice(alphabets)
        out.append(bytes(rnd.choice(a) for _ in range(rnd.randint(1, max_len))))
    return out


def test_device_partitioner_equals_the_binary64_oracle(lib):
    # FnvLua52 (integer form, rounding on the low word) against the oracle's real doubles and the
    # library's host function, inline keys and keys read from a byte space, every R of the configs
    keys = random_keys(60000, 1) + [b"the", b"local", b"end", b"function", b"=", b"--"]
    for k in keys:
        h = lib.th_fnv(k, len(k))
        assert h == O.fnv_lua52(k), k
        for R in (10, 15):
            p = lib.th_partition_of(k, len(k), R)
            assert p == h % R == _native.partition_of(k, R), (k, R)
    # SURVEY 8c known answers (R = 15)
    assert [lib.th_partition_of(k, len(k), 15) for k in (b"the", b"local", b"end", b"function", b"=", b"--")] == [8, 9, 4, 3, 4, 1]
    assert lib.th_fnv(b"the", 3) == 1093956953


def test_sort_words_give_the_reference_key_order(lib):
    # comparing the words of round 0, 1, 2, ... lexicographically is comparing the byte strings the
    # way Lua's `<` does in the C locale (unsigned bytes, the shorter of two equal prefixes first)
    keys = random_keys(4000, 2, max_len=30) + [b"a", b"a\x00", b"a\x00\x00", b"ab", b"abcdefg", b"abcdefgh", b"abcdefg\x00",
                                               b"abcdefghijklmno", b"abcdefghijklmnop", b"abcdefghijklmn\x00"]

    def words(k):
        return [lib.th_sort_word(k, len(k), r) for r in range((30 + 6) // 7 + 1)]
    w = {k: words(k) for k in set(keys)}
    ks = sorted(w)
    assert ks == sorted(w, key=lambda k: w[k])                      # same total order
    assert len({tuple(v) for v in w.values()}) == len(w)            # distinct keys, distinct word tuples
    for k in ks[:500]:                                              # the word itself: 7 bytes big-endian, min(rest, 8)
        for r in range(3):
            rest = k[7 * r:]
            want = (int.from_bytes(rest[:7].ljust(7, b"\0"), "big") << 8 | min(len(rest), 8)) if rest else 0
            assert w[k][r] == want, (k, r)


def test_owner_map_is_upper_bound_over_the_splitters(lib):
    rnd = random.Random(3)
    for world in (1, 2, 3, 8):
        assert lib.th_owner_of(0, None, None, world, 7, 123) == 7 % world          # no splitters: partition mod world
        sp = sorted((rnd.randrange(10), rnd.getrandbits(64)) for _ in range(world - 1))
        parts = (C.c_uint32 * max(1, len(sp)))(*[p for p, _ in sp])
        wordsv = (C.c_uint64 * max(1, len(sp)))(*[w for _, w in sp])
        probes = [(rnd.randrange(10), rnd.getrandbits(64)) for _ in range(300)] + sp
        for p, w in probes:
            got = lib.th_owner_of(len(sp), parts, wordsv, world, p, w)
            assert got == bisect.bisect_right(sp, (p, w)), (world, p, w)
