"""BASELINE config 4 (shape): bigram count -- many distinct keys, most longer than the 15-byte inline
key, so the key pool, the long-key compare and the tie rounds of the sort all get exercised."""
import random

import numpy as np
import pytest

from helpers import assert_rows_equal
from lua_mapreduce_1_b200 import _native, corpus
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def run_bigram(splits, R, max_keys=1 << 18, **kw):
    total = sum(len(s) for s in splits)
    ctx = _native.Context(R, total + 64 * len(splits) + 4096, max_keys=max_keys,
                          map_kind=_native.MR_MAP_BIGRAM_WS, **kw)
    try:
        for i, s in enumerate(splits):
            ctx.map_split(i, s)
        rows = ctx.rows()
        return rows, ctx.job_sync()
    finally:
        ctx.close()


def check(splits, R, **kw):
    want = O.run_bigram(splits, R)
    got, info = run_bigram(splits, R, **kw)
    assert info.n_tokens == want.tokens           # pairs emitted
    assert_rows_equal(got, want, f"bigram R={R}")
    return got, info


@pytest.mark.parametrize("text", [b"", b"one", b"a b", b"a b  c\td\r\ne f\n\ng\nh i", b"a\nb\nc", b"x  \n  y z ",
                                  b"a b a b a b\na b", b"p" * 9 + b" " + b"q" * 9, b"w " * 40])
def test_tiny_inputs(text):
    check([text], 10)


def test_pairs_never_span_lines_or_splits():
    got, _ = check([b"end of", b"split one\nnext line"], 10)
    keys = got.keys()
    assert b"of split" not in keys and b"one next" not in keys and b"end of" in keys and b"split one" in keys


def test_torture_and_long_keys():
    check([corpus.TORTURE], 15)
    big = b"Q" * 3000
    check([b"a " + big + b" b " + big + b"\n" + big + b" " + big + b" " + big], 10)


def test_many_keys_sharing_prefixes():
    rnd = random.Random(9)
    words = [b"international", b"internationally", b"internationalisation", b"inter", b"the", b"of", b"a"]
    words += [bytes(rnd.choice(b"abc") for _ in range(rnd.randrange(1, 9))) for _ in range(60)]
    lines = [b" ".join(rnd.choice(words) for _ in range(rnd.randrange(1, 30))) for _ in range(3000)]
    got, info = check([b"\n".join(lines[:1500]), b"\r\n".join(lines[1500:])], 10)
    assert info.tie_rounds >= 1


def test_synthetic_corpus_bigrams():
    splits, _ = corpus.generate(lines=[6000, 5000, 1, 3000])
    got, info = check(splits, 10, max_keys=1 << 20, table_slots=1 << 21)
    assert info.n_keys > 200_000                  # high cardinality: most pairs are distinct
    lens = np.diff(got.key_off.astype(np.int64))
    assert (lens > 15).sum() > 10_000             # the key pool really is used
