"""The device helpers of csrc/common.cuh and csrc/reduce_kernels.cuh, compiled for the host
(tests/host_helpers.cu): the partitioner, the sort words, the owner map, the tokeniser's whitespace
classifier and the radix sort's digit plan -- the source the kernels run -- checked on CPU against the
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
    deps = [SRC, os.path.join(ROOT, "lua-mapreduce-1_b200", "csrc", "common.cuh"),
            os.path.join(ROOT, "lua-mapreduce-1_b200", "csrc", "reduce_kernels.cuh"),   # map_kernel.cuh uses the same helpers
            os.path.join(ROOT, "lua-mapreduce-1_b200", "csrc", "msd_kernels.cuh")]
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
    L.th_space_flags.restype = C.c_uint32
    L.th_space_flags.argtypes = [C.c_uint32]
    L.th_nonspace_nibble.restype = C.c_uint32
    L.th_nonspace_nibble.argtypes = [C.c_uint32]
    L.th_digit_plan.restype = C.c_uint32
    L.th_digit_plan.argtypes = [C.c_uint64, C.c_uint32, C.POINTER(C.c_uint8), C.POINTER(C.c_uint32)]
    L.th_digit.restype = C.c_uint32
    L.th_digit.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint64]
    L.th_token_masks.restype = None
    L.th_token_masks.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint32)]
    L.th_token_len.restype = C.c_uint32
    L.th_token_len.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32]
    L.th_inline_key.restype = None
    L.th_inline_key.argtypes = [C.POINTER(C.c_uint32), C.c_uint32, C.c_uint32, C.POINTER(C.c_uint64)]
    L.th_scan_long_token.restype = C.c_uint32
    L.th_scan_long_token.argtypes = [C.c_char_p, C.c_uint64, C.c_uint32, C.POINTER(C.c_uint32)]
    L.th_same_long_key.restype = C.c_int
    L.th_same_long_key.argtypes = [C.c_char_p, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]
    L.th_op_roundtrip.restype = C.c_uint64
    L.th_op_roundtrip.argtypes = [C.c_uint32, C.c_uint64, C.c_uint64]
    L.th_msd_extract.restype = C.c_uint32
    L.th_msd_extract.argtypes = [C.POINTER(C.c_uint32), C.c_uint32, C.c_uint64, C.c_uint32, C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint32)]
    L.th_msd_pass_digit.restype = C.c_uint32
    L.th_msd_pass_digit.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32]
    for f in (L.th_inline_tail_less, L.th_rounds_less):
        f.restype = C.c_int
        f.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_int]
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


def test_whitespace_flags_are_isspace_in_every_byte_lane(lib):
    # mapfn.lua:5 splits on %s = C-locale isspace: 0x20 and 0x09..0x0D.  The SWAR classifier must be
    # exact in each byte lane whatever the neighbouring bytes are (no borrow from lane to lane)
    space = {0x20, 0x09, 0x0A, 0x0B, 0x0C, 0x0D}
    rnd = random.Random(4)
    edge = [0x00, 0x08, 0x09, 0x0D, 0x0E, 0x1F, 0x20, 0x21, 0x7F, 0x80, 0x88, 0x89, 0x8D, 0xA0, 0xFF]
    words = [a | b << 8 | c << 16 | d << 24 for a in edge for b in edge for c in edge for d in edge[::3]]
    words += [rnd.getrandbits(32) for _ in range(100000)]
    words += [b | (b2 << (8 * lane)) for b in range(256) for b2 in (0x09, 0x20, 0xFF) for lane in (1, 2, 3)]
    for x in words:
        want = sum(0x80 << (8 * i) for i in range(4) if (x >> (8 * i)) & 0xFF in space)
        assert lib.th_space_flags(x) == want, hex(x)
        nib = sum(1 << i for i in range(4) if (x >> (8 * i)) & 0xFF not in space)
        assert lib.th_nonspace_nibble(x) == nib, hex(x)


def test_digit_plan_covers_the_varying_bits_once_and_sorts_like_the_keys(lib):
    rnd = random.Random(5)
    MAXP = 12
    for trial in range(300):
        vlo = rnd.getrandbits(64) & rnd.getrandbits(64) if trial % 3 else rnd.getrandbits(64)
        vhi = rnd.getrandbits(32) & rnd.getrandbits(32) & (0xF if trial % 2 else 0xFFFFFFFF)
        if trial == 0:
            vlo, vhi = 0, 0
        if trial == 1:
            vlo, vhi = 2**64 - 1, 2**32 - 1
        out = (C.c_uint8 * (MAXP * 8 * 3))()
        nruns = (C.c_uint32 * MAXP)()
        np_ = lib.th_digit_plan(vlo, vhi, out, nruns)
        nbits = bin(vlo).count("1") + bin(vhi).count("1")
        assert np_ == (nbits + 7) // 8
        covered = 0
        for p in range(np_):
            used = 0
            assert 1 <= nruns[p] <= 8
            for r in range(nruns[p]):
                shift, width, dst = out[(p * 8 + r) * 3], out[(p * 8 + r) * 3 + 1], out[(p * 8 + r) * 3 + 2]
                assert dst == used and width >= 1 and not (shift < 64 < shift + width)    # no run straddles lo/hi
                mask = ((1 << width) - 1) << shift
                assert mask & covered == 0 and mask & (vlo | vhi << 64) == mask           # only varying bits, each once
                assert mask > covered or covered == 0                                     # least significant first
                covered |= mask
                used += width
            assert used == 8 or p == np_ - 1
        assert covered == vlo | vhi << 64
        # sorting by the digits, least significant pass first (stable), is sorting by the varying bits
        if nbits and trial < 60:
            keys = [(rnd.getrandbits(32), rnd.getrandbits(64)) for _ in range(200)]
            order = list(range(len(keys)))
            for p in range(np_):
                order.sort(key=lambda i: lib.th_digit(vlo, vhi, p, keys[i][0], keys[i][1]))
            want = sorted(range(len(keys)), key=lambda i: ((keys[i][0] & vhi), (keys[i][1] & vlo), ))
            got_keys = [((keys[i][0] & vhi), (keys[i][1] & vlo)) for i in order]
            assert got_keys == [((keys[i][0] & vhi), (keys[i][1] & vlo)) for i in want]


def test_token_bounds_from_masks_are_what_splitting_on_whitespace_gives(lib):
    # a lane of the map kernel sees its 32 bytes, the 32 that follow and whether the byte before is a
    # token byte; where tokens start, how long the short ones are and which class each falls in must
    # be what bytes.split() says about the same 65 bytes
    rnd = random.Random(6)
    space = b" \t\n\r\x0b\x0c"
    for trial in range(3000):
        p_space = rnd.choice([0.05, 0.15, 0.3, 0.6, 0.95])
        data = bytes(rnd.choice(space) if rnd.random() < p_space else rnd.randrange(0x21, 0x100) for _ in range(65))
        if trial % 7 == 0:                                   # long runs: tokens of 16+ bytes, tokens that leave the window
            data = bytes(0x41 if (i // rnd.choice([9, 17, 33])) % 2 else 0x20 for i in range(65))
        carry, body = int(data[0] not in space), data[1:]
        nib = lambda b4: lib.th_nonspace_nibble(int.from_bytes(b4, "little"))
        m = sum(nib(body[4 * i:4 * i + 4]) << (4 * i) for i in range(8))
        nxt = sum(nib(body[32 + 4 * i:36 + 4 * i]) << (4 * i) for i in range(8))
        out = (C.c_uint32 * 4)()
        lib.th_token_masks(m, nxt, carry, out)
        starts, s_long, s_med, s_short = out
        for i in range(32):
            is_start = body[i] not in space and (data[i] in space)          # data[i] is the byte before body[i]
            assert (starts >> i) & 1 == int(is_start), (trial, i)
            if not is_start:
                assert not ((s_long | s_med | s_short) >> i) & 1
                continue
            n = 0
            while i + n < 64 and body[i + n] not in space:
                n += 1
            cls = "long" if n >= 16 else "med" if n >= 8 else "short"       # 16+ seen in the window is 16+ for sure
            assert ((s_long >> i) & 1, (s_med >> i) & 1, (s_short >> i) & 1) == \
                (int(cls == "long"), int(cls == "med"), int(cls == "short")), (trial, i, n)
            if cls != "long":
                assert lib.th_token_len(m, nxt, i) == n, (trial, i, n)


def test_inline_keys_built_from_staged_words_hold_the_token_bytes_and_length(lib):
    # a token at any byte position of the staging buffer: the short key (<= 7 bytes) is its bytes
    # little-endian with the length in the top byte of ONE 64-bit word; the medium key (8..15 bytes)
    # is bytes 0..7 in k0, bytes 8..14 and the length << 56 in k1 -- never a byte of the neighbours
    rnd = random.Random(7)
    for trial in range(4000):
        buf = bytes(rnd.randrange(1, 256) for _ in range(40))
        words = (C.c_uint32 * 10)(*[int.from_bytes(buf[4 * i:4 * i + 4], "little") for i in range(10)])
        pos, n = rnd.randrange(0, 20), rnd.randint(1, 15)
        tok = buf[pos:pos + n]
        out = (C.c_uint64 * 2)()
        lib.th_inline_key(words, pos, n, out)
        if n <= 7:
            assert out[0] == int.from_bytes(tok, "little") | n << 56 and out[1] == 0, (pos, n)
        else:
            assert out[0] == int.from_bytes(tok[:8], "little"), (pos, n)
            assert out[1] == int.from_bytes(tok[8:], "little") | n << 56, (pos, n)


def test_long_token_scanner_and_long_key_equality(lib):
    rnd = random.Random(8)
    space = b" \t\n\r\x0b\x0c"
    for trial in range(1500):
        n1, n2 = rnd.randint(1, 300), rnd.randint(1, 300)
        t1 = bytes(rnd.choice([rnd.randrange(0x21, 0x100), 0x00, 0x0E, 0x1F]) for _ in range(n1))
        t2 = t1 if trial % 3 == 0 else bytes(rnd.choice(t1) for _ in range(n2))
        pad = rnd.randrange(0, 9)
        buf = b"x" * pad + t1 + bytes([rnd.choice(space)]) + t2 + bytes([rnd.choice(space)]) + b"tail" + b" " * 32
        a, b = pad, pad + len(t1) + 1
        h1, h2 = C.c_uint32(), C.c_uint32()
        assert lib.th_scan_long_token(buf, a, 1 << 20, C.byref(h1)) == len(t1)       # stops at the first whitespace byte
        assert lib.th_scan_long_token(buf, b, 1 << 20, C.byref(h2)) == len(t2)
        assert lib.th_scan_long_token(buf, a, len(t1) - 1, C.byref(C.c_uint32())) == 0 or len(t1) == 0   # too long: refused
        if t1 == t2:
            assert h1.value == h2.value                                              # the hash depends on the bytes only
        same = lib.th_same_long_key(buf, a, b, len(t1), len(t2), h1.value, h2.value)
        assert same == int(t1 == t2), (trial, t1, t2)


def test_reducer_encodings_have_identity_zero_and_merge_correctly(lib):
    rnd = random.Random(9)
    SUM, MIN, MAX, COUNT = 0, 1, 2, 3
    for _ in range(2000):
        a, b = rnd.getrandbits(rnd.choice([1, 8, 32, 52])), rnd.getrandbits(rnd.choice([1, 8, 32, 52]))
        assert lib.th_op_roundtrip(SUM, a, b) == a + b
        assert lib.th_op_roundtrip(MIN, a, b) == min(a, b)
        assert lib.th_op_roundtrip(MAX, a, b) == max(a, b)
        assert lib.th_op_roundtrip(COUNT, a, b) == 2


def test_msd_plan_bins_and_low_words_order_like_the_keys_and_unit_passes_sort_them(lib):
    # msd_kernels.cuh: (bin, packed low word) must order the sort words exactly like (partition, key
    # bytes, length) does; the shared-memory passes of a unit, least significant first and stable, must
    # order (local bin, low) numerically
    from helpers import sort_word
    rnd = random.Random(17)
    alphabets = [b"abcdefghijklmnopqrstuvwxyz", b"ab", bytes(range(256)), b"0123456789", b"xyz_./-ABC\x00\xff"]
    for trial in range(120):
        alpha = alphabets[trial % len(alphabets)]
        R = rnd.choice([1, 7, 10, 15, 31, 1000, 1 << 24])
        nkeys = rnd.choice([1, 2, 40, 300])
        keys = {bytes(rnd.choice(alpha) for _ in range(rnd.choice([1, 1, 2, 3, 5, 7, 8, 9, 12, 20]))) for _ in range(nkeys)}
        words = sorted({(rnd.randrange(R), sort_word(k)) for k in keys})
        occ = (C.c_uint32 * 56)()
        for _, w in words:
            for p in range(7):
                b = (w >> (56 - 8 * p)) & 0xFF
                occ[p * 8 + (b >> 5)] |= 1 << (b & 31)
        n = rnd.choice([len(words), 5000, 524219, 3_000_000, 1 << 30])
        out = (C.c_uint32 * 5)()
        low = C.c_uint64()
        packed = []
        for hi, lo in words:
            d = lib.th_msd_extract(occ, R, n, hi, lo, C.byref(low), out)
            packed.append((d, low.value))
        nbins, k, e, low_bits, fallback = list(out)
        assert 1 <= nbins <= max(1, min(n // 8, 1 << 15)) and 0 <= k <= 9 and e >= 1
        if fallback:
            assert low_bits > 64
            continue
        assert all(d < nbins and (low_bits == 64 or l < (1 << low_bits)) for d, l in packed)
        assert packed == sorted(packed) and len(set(packed)) == len(words)          # same order, nothing merged
        dmin = min(d for d, _ in packed)
        items = [(d - dmin, l) for d, l in packed]
        rnd.shuffle(items)
        lbits = max(b for b, _ in items).bit_length()
        npass = (low_bits + lbits + 7) // 8
        order = list(range(len(items)))
        for p in range(npass):
            order.sort(key=lambda i: lib.th_msd_pass_digit(items[i][1], items[i][0], low_bits, p))
        assert [items[i] for i in order] == sorted(items)


def inline_key(b):
    """(k0, k1) of a key of 1..15 bytes as the kernels hold it: bytes 0..7 little-endian in k0, bytes
    8..14 in k1's low 56 bits, the length in k1's top byte (common.cuh)."""
    assert 1 <= len(b) <= 15
    padded = b + bytes(16 - len(b))
    return int.from_bytes(padded[:8], "little"), int.from_bytes(padded[8:15], "little") | len(b) << 56


def test_one_word_order_of_inline_keys_that_share_seven_bytes(lib):
    # the tie groups of msd_unit_kernel: keys equal in bytes 0..6, 8..15 bytes long.  The one-compare
    # order must be the byte order with the shorter key first (Python's bytes order), and must agree with
    # the round-by-round order key_less takes from sort_chunk -- NUL bytes, prefixes, lengths 14/15 included
    rnd = random.Random(12)
    prefix = b"PREFIX_"
    tails = [b"\x00", b"\x00\x00", b"a", b"a\x00", b"a\x00b", b"ab", b"\xff", b"\xff\x00", b"1234567", b"12345670",
             b"1234567\x00", b"1234568", b"12345678", b"\x00" * 8, b"\x00" * 7, b"\x00" * 7 + b"\x01"]
    tails += [bytes(rnd.choice([0, 0, 1, 97, 98, 255]) for _ in range(rnd.randrange(1, 9))) for _ in range(300)]
    keys = sorted({prefix + t for t in tails})
    for a in keys:
        for b in rnd.sample(keys, 40):
            ka, kb = inline_key(a), inline_key(b)
            for first in (0, 1):
                want = a < b or (a == b and bool(first))
                assert bool(lib.th_inline_tail_less(*ka, *kb, first)) == want, (a, b, first)
                assert bool(lib.th_rounds_less(*ka, *kb, first)) == want, (a, b, first)
