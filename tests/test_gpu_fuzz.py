"""Randomised differential parity: generated texts of many shapes (tiny alphabets = long common
prefixes and heavy ties, raw bytes, every whitespace byte, tokens from 1 to 700 bytes, empty and
one-byte splits, R from 1 to 31) through the C ABI against the oracle, bit-exact.  Fixed seeds: a
failure names its case and can be replayed alone."""
import os
import random

import numpy as np
import pytest

from helpers import assert_rows_equal
from lua_mapreduce_1_b200 import _native
from oracle import oracle as O

pytestmark = pytest.mark.gpu
EXTRA = int(os.environ.get("MRGPU_FUZZ_EXTRA", "0"))     # more seeds for a one-off deeper run

SPACES = [b" ", b"\t", b"\n", b"\r", b"\x0b", b"\x0c", b"\r\n", b"  ", b" \n ", b"\n\n"]
ALPHABETS = {
    "ab": b"ab",                                   # two letters: almost every key ties with another
    "lower": bytes(range(0x61, 0x7b)),
    "printable": bytes(range(0x21, 0x7f)),
    "bytes": bytes(b for b in range(256) if b not in (0x20, 0x09, 0x0a, 0x0b, 0x0c, 0x0d)),   # NUL and >= 0x80 too
    "digits/": b"0123456789/",
}


def make_case(seed):
    rnd = random.Random(seed)
    alpha = ALPHABETS[rnd.choice(sorted(ALPHABETS))]
    vocab_n = rnd.choice([1, 3, 40, 500, 5000])
    max_len = rnd.choice([1, 3, 7, 8, 15, 16, 17, 40, 700])
    prefix = bytes(rnd.choice(alpha) for _ in range(rnd.choice([0, 0, 6, 7, 8, 14, 20])))

    def word():
        n = rnd.randint(1, max_len)
        return prefix[:max(0, max_len - 1)][:rnd.randint(0, len(prefix))] + bytes(rnd.choice(alpha) for _ in range(n)) \
            if prefix else bytes(rnd.choice(alpha) for _ in range(n))
    vocab = [word() for _ in range(vocab_n)]
    nsplits = rnd.choice([1, 1, 2, 5, 20])
    splits = []
    for _ in range(nsplits):
        ntok = rnd.choice([0, 0, 1, 2, 50, 2000])
        parts = [rnd.choice(SPACES)] if rnd.random() < 0.3 else []
        for _ in range(ntok):
            parts.append(rnd.choice(vocab) if rnd.random() < 0.9 else word())
            parts.append(rnd.choice(SPACES))
        if parts and rnd.random() < 0.5:
            parts.pop()                               # no trailing whitespace
        splits.append(b"".join(parts))
    return splits, rnd.choice([1, 2, 7, 10, 15, 16, 31])


def total_bytes(splits):
    return sum(len(s) for s in splits)


@pytest.mark.parametrize("seed", list(range(40)) + list(range(1000, 1000 + EXTRA)))
def test_word_count_matches_the_oracle(seed):
    splits, R = make_case(seed)
    want = O.run_naive(splits, R)
    ctx = _native.Context(R, total_bytes(splits) + 64 * len(splits) + 4096, max_keys=1 << 16)
    try:
        for i, s in enumerate(splits):
            ctx.map_split(i, s)
        ctx.reduce_all()
        info = ctx.job_sync()
        assert info.n_tokens == want.tokens, f"seed {seed}"
        assert_rows_equal(ctx.rows(), want, f"seed {seed} R={R}")
        if seed % 4 == 0:
            # the same job through the peer-to-peer shuffle against the context's own window
            ctx.p2p_attach([ctx.p2p_setup(max(4096, 64 * (total_bytes(splits) + 64) // 32 * 32))])
            ctx.job_begin(keep_input=True)
            ctx.map_finish()
            ctx.shuffle_p2p()
            ctx.reduce_all()
            assert_rows_equal(ctx.rows(), want, f"seed {seed} R={R} through p2p")
    finally:
        ctx.close()


@pytest.mark.parametrize("seed", list(range(100, 116)) + list(range(5000, 5000 + EXTRA // 4)))
def test_bigram_count_matches_the_oracle(seed):
    splits, R = make_case(seed)
    splits = [s for s in splits if max((len(t) for t in s.split()), default=0) <= 300]   # pairs of 700-byte tokens: out of the pool's range here
    want = O.run_bigram(splits, R)
    ctx = _native.Context(R, total_bytes(splits) + 64 * len(splits) + 4096, max_keys=1 << 16,
                          map_kind=_native.MR_MAP_BIGRAM_WS)
    try:
        for i, s in enumerate(splits):
            ctx.map_split(i, s)
        ctx.reduce_all()
        info = ctx.job_sync()
        assert info.n_tokens == want.tokens, f"seed {seed}"
        assert_rows_equal(ctx.rows(), want, f"bigram seed {seed} R={R}")
    finally:
        ctx.close()


@pytest.mark.parametrize("seed", range(200, 208))
def test_number_keys_match_the_oracle(seed):
    rng = np.random.default_rng(seed)
    n = int(rng.choice([0, 1, 17, 5000, 200000]))
    space = int(rng.choice([1, 7, 1000, 2**32]))
    batches = []
    for _ in range(int(rng.integers(1, 4))):
        k = rng.integers(0, space, n, dtype=np.uint64).astype(np.uint32)
        if n:
            k[rng.integers(0, n, max(1, n // 3))] = np.uint32(space - 1)       # a heavy hitter
        batches.append((k, rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32)))
    R = int(rng.choice([1, 3, 10, 16]))
    for kind, op in ((_native.MR_REDUCE_SUM_U64, 0), (_native.MR_REDUCE_MIN_U64, 1), (_native.MR_REDUCE_MAX_U64, 2),
                     (_native.MR_REDUCE_COUNT, 3)):
        want = O.groupby_u32(batches, R, op)
        ctx = _native.Context(R, 1 << 23, max_keys=1 << 18, map_kind=_native.MR_MAP_RECORDS_U32,
                              partition_kind=_native.MR_PARTITION_MOD_U32, reduce_kind=kind)
        try:
            for i, (k, v) in enumerate(batches):
                ctx.map_records(i, k, v)
            ctx.reduce_all()
            got = ctx.rows()
            assert got.n == want.n, (seed, kind)
            assert np.array_equal(got.part, want.part) and np.array_equal(got.count, want.count), (seed, kind)
            assert np.array_equal(got.keys_u32(), np.frombuffer(want.key_bytes.tobytes(), ">u4")), (seed, kind)
        finally:
            ctx.close()
