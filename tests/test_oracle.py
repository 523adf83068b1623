"""The CPU oracle against the reference's own known answers -- no GPU needed."""
import random

import numpy as np

from helpers import golden, rows_digest
from oracle import oracle as O


def test_isspace_is_c_locale():
    ws = {0x20, 0x09, 0x0A, 0x0B, 0x0C, 0x0D}
    for c in range(256):
        assert bool(O.lib().mro_is_space(c)) == (c in ws)


def test_partition_hash_known_answers():
    # double-FNV hash and partition with R=15 / R=10
    kat = {b"the": (1093956953, 8, 3), b"local": (2884676004, 9, 4), b"end": (1091544064, 4, 4),
           b"function": (904614438, 3, 8), b"=": (84696349, 4, 9), b"--": (4285954906, 1, 6)}
    for k, (h, p15, p10) in kat.items():
        assert O.fnv_lua52(k) == h
        assert O.partition(k, 15) == p15 and O.partition(k, 10) == p10
    # it is NOT uint32 FNV-1
    h = 2166136261
    for c in b"the":
        h = (h * 16777619) % 2**32 ^ c
    assert h == 996983870 and h != O.fnv_lua52(b"the")


def test_partition_integer_emulation_matches_double():
    rnd = random.Random(7)
    for _ in range(50000):
        k = bytes(rnd.randrange(256) for _ in range(rnd.randrange(0, 24)))
        assert O.fnv_lua52(k) == O.fnv_lua52_int(k)


def test_heap_kat():
    # heap.lua:99-118
    assert O.heap_sort([20, 10, 15, 1]) == [1, 10, 15, 20]
    assert O.heap_sort([]) == []
    assert O.heap_sort([30]) == [30]


def test_merge_iterator_kat():
    # utils.lua:360-380: f1 = {1:{1,1}, 2:{1}, 3:{1}}, f2 = {1:{1,1,1,1}, 3:{1}, 4:{1}}
    got = O.merge_kat([[(1, 2), (2, 1), (3, 1)], [(1, 4), (3, 1), (4, 1)]])
    assert got == [(1, 6), (2, 1), (3, 2), (4, 1)]


def test_key_order_is_unsigned_bytes_shorter_first():
    # utils.lua:349 keys_sorted{c,a,b} and Lua string `<`
    keys = [b"c", b"a", b"b", b"ab", b"a\x00", b"\xff", b"\x7f", b"abc", b"ab\xc3\xa9"]
    want = sorted(keys)
    r = O.run_naive([b" ".join(keys)], 1)
    assert r.keys() == want
    assert O.key_cmp(b"a", b"ab") < 0 and O.key_cmp(b"\x80", b"\x7f") > 0 and O.key_cmp(b"x", b"x") == 0


def test_fixture_known_answers(reference_fixture):
    g = golden("wordcount_fixture.json")
    assert sum(map(len, reference_fixture)) == g["bytes"] == 47469
    for R in (15, 10):
        naive = O.run_naive(reference_fixture, R)
        staged = O.run_staged(reference_fixture, R, 4)
        assert naive.same_as(staged)
        assert naive.tokens == g["tokens"] == 4977 and naive.n == g["distinct"] == 1575
        d = naive.as_dict()
        for k, v in g["top"]:
            assert d[k.encode()] == v
        assert np.bincount(naive.part, minlength=R).tolist() == g[f"R{R}"]["keys_per_partition"]
        assert rows_digest(naive.part, naive.count, naive.key_off, naive.key_bytes) == g[f"R{R}"]["sha256"]
    assert g["R15"]["keys_per_partition"] == [106, 103, 116, 121, 102, 83, 121, 98, 92, 107, 116, 103, 104, 95, 108]


def test_torture_rows_match_golden():
    from lua_mapreduce_1_b200 import corpus
    g = golden("torture_rows.json")
    for R in (15, 10, 1):
        for r in (O.run_naive([corpus.TORTURE], R), O.run_staged([corpus.TORTURE], R, 2)):
            want = g[f"R{R}"]["rows"]
            assert r.n == len(want)
            assert [[int(p), k.hex(), int(c)] for p, k, c in zip(r.part, r.keys(), r.count)] == want
            assert rows_digest(r.part, r.count, r.key_off, r.key_bytes) == g[f"R{R}"]["sha256"]


def test_staged_equals_naive_on_ragged_splits():
    rnd = random.Random(3)
    words = [bytes(rnd.choice(b"abcdefgh") for _ in range(rnd.randrange(1, 12))) for _ in range(300)]
    splits = []
    for s in range(7):
        n = [0, 1, 5, 400, 3000, 17, 0][s]
        splits.append(b"".join(rnd.choice(words) + rnd.choice([b" ", b"\n", b"\t ", b"  "]) for _ in range(n)))
    splits.append(b"")
    splits.append(b"   \n\n")
    splits.append(b"last-token-without-newline")
    a, b = O.run_naive(splits, 10), O.run_staged(splits, 10, 3)
    assert a.same_as(b) and a.tokens == b.tokens
    assert int(a.count.sum()) == a.tokens


def test_groupby_oracle_on_number_keys():
    # numbers order numerically (Lua `<`), partition = key % R, values summed (reducefn.lua:1-5)
    k = np.array([5, 3, 5, 0, 4294967295, 3, 5, 10], dtype=np.uint32)
    v = np.array([1, 2, 3, 4, 5, 6, 7, 8], dtype=np.uint32)
    r = O.groupby_u32([(k[:4], v[:4]), (k[4:], v[4:])], 10)
    keys = np.frombuffer(r.key_bytes.tobytes(), ">u4")
    assert r.part.tolist() == [0, 0, 3, 5, 5] and keys.tolist() == [0, 10, 3, 5, 4294967295]
    assert r.count.tolist() == [4, 8, 8, 11, 5] and r.tokens == 8
    from lua_mapreduce_1_b200.examples import GroupBySum
    assert [GroupBySum.partitionfn(int(x)) for x in keys] == r.part.tolist()


def test_bigram_oracle_against_python_restatement():
    import collections
    from lua_mapreduce_1_b200 import corpus
    splits = [b"a b  c\td\r\ne f\n\ng\nh i", b"x y", corpus.TORTURE, corpus.generate_split(3, 300)[0].tobytes()]
    want = collections.Counter()
    for s in splits:
        for line in s.split(b"\n"):              # io.lines cuts at "\n" only
            w = line.split()
            for a, b in zip(w, w[1:]):
                want[a + b" " + b] += 1
    r = O.run_bigram(splits, 10)
    assert r.as_dict() == dict(want) and r.tokens == sum(want.values())
    assert r.keys() == [k for _, k in sorted((O.partition(k, 10), k) for k in want)]


def test_low_word_rounding_used_on_the_device_equals_the_double_product():
    # common.cuh FnvLua52::byte rounds the 57-bit product h * 16777619 to 53 bits on its low word
    # only; here the same formula in numpy against real doubles, two million states h incl. ties
    import numpy as np
    rng = np.random.default_rng(5)
    h = np.concatenate([rng.integers(0, 2**32, 2_000_000, dtype=np.uint64),
                        np.arange(0xFFFF0000, 2**32, dtype=np.uint64), np.arange(0, 70000, dtype=np.uint64)])
    p = h.astype(np.float64) * 16777619.0                                   # IEEE product, rounded to 53 bits
    want = (p - np.floor(p / 4294967296.0) * 4294967296.0).astype(np.uint64)
    full = h * np.uint64(16777619)
    lo, hi = full & np.uint64(0xFFFFFFFF), full >> np.uint64(32)
    nbits = np.zeros(len(h), dtype=np.int64)
    nz = hi > 0
    nbits[nz] = np.floor(np.log2(hi[nz].astype(np.float64))).astype(np.int64) + 1   # 32 - clz(hi)
    s = np.maximum(nbits - 21, 0).astype(np.uint64)                        # 11 - clz(hi)
    sm1 = np.where(s > 0, s - np.uint64(1), np.uint64(0))
    add = np.where(s > 0, (np.uint64(1) << sm1) - np.uint64(1) + ((lo >> s) & np.uint64(1)), np.uint64(0))
    got = (((lo + add) >> s) << s) & np.uint64(0xFFFFFFFF)
    assert np.array_equal(got, want)
