"""The two reduce sides -- units of bins sorted in shared memory (msd_kernels.cuh) and the global LSD
radix sort with its tie machinery (reduce_kernels.cuh) -- must give the same rows, which are the
oracle's.  mr_job_info.flags says which one produced them; MRGPU_REDUCE=lsd forces the second."""
import os
import random

import numpy as np
import pytest

from helpers import assert_rows_equal
from lua_mapreduce_1_b200 import _native, corpus
from oracle import oracle as O

pytestmark = pytest.mark.gpu
LSD, DECLINED = 1, 2


def run(splits, R, force_lsd, max_keys=1 << 18, jobs=1, **kw):
    old = os.environ.pop("MRGPU_REDUCE", None)
    if force_lsd:
        os.environ["MRGPU_REDUCE"] = "lsd"
    try:
        total = sum(len(s) for s in splits)
        ctx = _native.Context(R, total + 64 * len(splits) + 4096, max_keys=max_keys, **kw)
    finally:
        os.environ.pop("MRGPU_REDUCE", None)
        if old is not None:
            os.environ["MRGPU_REDUCE"] = old
    try:
        for i, s in enumerate(splits):
            ctx.map_split(i, s)
        out = []
        for j in range(jobs):
            if j:
                ctx.job_begin(keep_input=True)
            ctx.map_finish()
            ctx.reduce_all()
            info = ctx.job_sync()
            out.append((ctx.rows(), int(info.flags), int(info.tie_rounds)))
        return out
    finally:
        ctx.close()


def both(splits, R, expect_msd=True, **kw):
    want = O.run_naive(splits, R)
    (a, fa, _), (a2, fa2, _) = run(splits, R, False, jobs=2, **kw)
    (b, fb, _), = run(splits, R, True, **kw)
    assert_rows_equal(a, want, "msd path")
    assert_rows_equal(a2, want, "msd path, second job")
    assert_rows_equal(b, want, "lsd path")
    assert np.array_equal(a.part_begin, b.part_begin)
    assert fb == LSD
    if expect_msd:
        assert fa == 0 and fa2 == 0, f"the shared-memory path declined: flags {fa:#x}"
    else:
        assert fa & DECLINED and fa2 == LSD      # declined once: the next jobs go straight to the LSD path
    return fa


def test_synthetic_text_takes_the_shared_memory_path():
    splits, _ = corpus.generate(lines=[3000, 2500, 1, 0, 1800])
    both(splits, 10)
    both(splits, 15)
    both(splits[:1], 1)


def test_tiny_and_empty_jobs():
    for text in (b"", b"a", b"b a", b"x" * 9 + b" " + b"x" * 8 + b" " + b"x" * 7, b"same same same"):
        both([text], 10)
    both([], 7)


def test_torture_split_and_long_keys():
    both([corpus.TORTURE, corpus.TORTURE[::-1]], 10)
    rnd = random.Random(3)
    words = [bytes(rnd.choice(b"abcdef") for _ in range(rnd.randrange(16, 90))) for _ in range(3000)]
    both([b" ".join(rnd.choice(words) for _ in range(20000))], 10)


def test_small_tie_groups_are_ordered_inside_the_unit():
    # keys that agree in their first 7 bytes, in groups of at most 16: ordered by the unit kernel itself
    rnd = random.Random(8)
    keys = []
    for g in range(400):
        prefix = bytes(rnd.choice(b"abcdefghij") for _ in range(7))
        keys += [prefix + bytes(rnd.choice(b"xyz") for _ in range(rnd.randrange(1, 9))) for _ in range(rnd.randrange(1, 12))]
    text = b" ".join(rnd.choice(keys) for _ in range(30000))
    (rows, flags, tie_rounds), = run([text], 10, False)
    assert flags == 0 and tie_rounds == 1
    assert_rows_equal(rows, O.run_naive([text], 10), "small tie groups")


def test_tie_groups_with_nul_bytes_prefixes_and_the_inline_limit():
    # one 7-byte prefix, then every awkward tail: none, NUL bytes (a zero-padded compare must still put
    # the shorter key first), one key a prefix of the next, tails that end exactly at 14 and 15 bytes
    # (the last inline length) and a 16+-byte key (kept by reference) -- all in one group, both paths
    rnd = random.Random(5)
    tails = [b"", b"\x00", b"\x00\x00", b"a", b"a\x00", b"a\x00b", b"ab", b"abc", b"b", b"\xff", b"\xff\x00",
             b"1234567", b"12345670", b"1234567\x00", b"1234568", b"12345678", b"123456789abcdef", b"z" * 30]
    keys = [b"PREFIX_" + t for t in tails] + [b"prefix2" + bytes([c]) * n for c in (0, 1, 255) for n in range(1, 9)]
    text = b" ".join(rnd.choice(keys) for _ in range(6000))
    both([text], 10)
    both([text, text[::-1]], 3)
    # the same keys arriving in copies (what an owner sees after a shuffle): two jobs' worth in one split list
    both([text] * 3, 10)


def test_large_tie_groups_make_the_shared_memory_path_decline():
    # ~250 keys per partition agree in their first 7 bytes: more than a unit orders by itself (64)
    rnd = random.Random(2)
    keys = sorted({b"prefix_" + bytes(rnd.choice(b"abcde") for _ in range(rnd.randrange(1, 12))) for _ in range(2500)})
    text = b" ".join(rnd.choice(keys) for _ in range(12000))
    fa = both([text], 10, expect_msd=False)
    assert (fa >> 8) & (8 | 4)                      # MSD_FB_TIES (or a bin too large for a unit)
    # up to 64 of them are the unit's own business
    keys = keys[:300]
    text = b" ".join(rnd.choice(keys) for _ in range(3000))
    both([text], 10, expect_msd=True)


def test_binary_keys_of_every_byte_value():
    rnd = random.Random(21)
    ws = b" \t\n\x0b\x0c\r"
    toks = [bytes(rnd.choice([c for c in range(256) if c not in ws]) for _ in range(rnd.randrange(1, 12))) for _ in range(5000)]
    both([b" ".join(rnd.choice(toks) for _ in range(40000))], 31)


def test_number_keys_group_by():
    rng = np.random.default_rng(4)
    n = 400_000
    keys = (rng.zipf(1.3, n) % 50_000).astype(np.uint32) * np.uint32(2654435761)
    vals = rng.integers(0, 100, n, dtype=np.uint32)
    want = O.groupby_u32([(keys, vals)], 10)
    for force in (False, True):
        if force:
            os.environ["MRGPU_REDUCE"] = "lsd"
        try:
            ctx = _native.Context(10, 8 * n + (1 << 16), max_keys=1 << 18, map_kind=_native.MR_MAP_RECORDS_U32,
                                  partition_kind=_native.MR_PARTITION_MOD_U32)
        finally:
            os.environ.pop("MRGPU_REDUCE", None)
        try:
            ctx.map_records(0, keys, vals)
            got = ctx.rows()
            flags = int(ctx.job_sync().flags)
        finally:
            ctx.close()
        assert_rows_equal(got, want, "group-by")
        assert flags == (LSD if force else 0)


def test_units_numbered_in_advance_or_by_look_back_give_the_same_rows():
    # a job without a shuffle has only distinct keys: the scatter kernel numbers every unit's rows and bytes
    # in advance; MRGPU_MSD_LOOKBACK makes the units find their numbers by look-back, as jobs after a shuffle do
    splits, _ = corpus.generate(lines=[4000, 3500, 1, 0, 2800])
    splits.append(corpus.TORTURE)
    want = O.run_naive(splits, 10)
    for lookback in (False, True):
        if lookback:
            os.environ["MRGPU_MSD_LOOKBACK"] = "1"
        try:
            (rows, flags, _), (rows2, flags2, _) = run(splits, 10, False, jobs=2)
        finally:
            os.environ.pop("MRGPU_MSD_LOOKBACK", None)
        assert flags == 0 and flags2 == 0
        assert_rows_equal(rows, want, f"lookback={lookback}")
        assert_rows_equal(rows2, want, f"lookback={lookback}, second job")
