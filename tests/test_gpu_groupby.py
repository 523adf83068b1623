"""BASELINE config 5 (shape): group-by-sum on integer keys, heavy hitters, uneven partitions --
bit-exact against the oracle through the C ABI."""
import numpy as np
import pytest

from helpers import assert_rows_equal
from lua_mapreduce_1_b200 import _native
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def ctx_for(R, n_records, **kw):
    return _native.Context(R, 8 * n_records + (1 << 16), max_keys=kw.pop("max_keys", 1 << 20),
                           map_kind=_native.MR_MAP_RECORDS_U32, partition_kind=_native.MR_PARTITION_MOD_U32, **kw)


def skewed(n, seed, distinct=200_000):
    rng = np.random.default_rng(seed)
    keys = (rng.zipf(1.2, n) % distinct).astype(np.uint32) * np.uint32(2654435761)   # spread over the u32 range
    hot = rng.random(n)
    keys[hot < 0.08] = 7                      # heavy hitters: > 5 % of the records each
    keys[(hot >= 0.08) & (hot < 0.14)] = 0
    keys[(hot >= 0.14) & (hot < 0.20)] = 0xFFFFFFFF
    vals = rng.integers(0, 1000, n, dtype=np.uint32)
    return keys, vals


@pytest.mark.parametrize("R", [10, 7, 1])
def test_small_and_edge_keys(R):
    k = np.array([5, 3, 5, 0, 4294967295, 3, 5, 10, 2**31, 2**31 - 1, 1 << 24, 255, 256], dtype=np.uint32)
    v = np.array([1, 2, 3, 4, 4294967295, 6, 7, 8, 4294967295, 1, 2, 3, 4], dtype=np.uint32)
    ctx = ctx_for(R, len(k))
    try:
        ctx.map_records(0, k[:6], v[:6])
        ctx.map_records(1, k[6:], v[6:])
        ctx.map_records(2, k[:0], v[:0])
        got = ctx.rows()
        info = ctx.job_sync()
    finally:
        ctx.close()
    want = O.groupby_u32([(k, v)], R)
    assert info.n_tokens == len(k)
    assert_rows_equal(got, want, f"groupby R={R}")
    assert np.array_equal(got.keys_u32() % R, got.part)
    # numeric order inside every partition
    for p in range(R):
        ks = got.keys_u32()[int(got.part_begin[p]):int(got.part_begin[p + 1])].astype(np.int64)
        assert np.all(np.diff(ks) > 0)


def test_skewed_group_by_sum_with_batches_and_keep_input():
    n = 3_000_000
    keys, vals = skewed(n, 5)
    want = O.groupby_u32([(keys, vals)], 10)
    ctx = ctx_for(10, n)
    try:
        for i, (a, b) in enumerate(zip(np.array_split(keys, 7), np.array_split(vals, 7))):
            ctx.map_records(i, a, b)
        ctx.map_finish()
        ctx.reduce_all()
        info = ctx.job_sync()
        got = ctx.rows()
        assert info.n_tokens == n and int(got.count.sum()) == int(vals.astype(np.uint64).sum())
        assert_rows_equal(got, want, "skewed group-by")
        sizes = np.diff(got.part_begin.astype(np.int64))
        assert sizes.sum() == got.n
        ctx.job_begin(keep_input=True)           # "loop" iteration over the resident records
        assert_rows_equal(ctx.rows(), want, "keep_input")
        ctx.job_begin()
        ctx.map_records(0, keys[:1000], vals[:1000])
        assert_rows_equal(ctx.rows(), O.groupby_u32([(keys[:1000], vals[:1000])], 10), "fresh job")
    finally:
        ctx.close()


def test_wrong_entry_point_is_refused():
    ctx = ctx_for(10, 1000)
    try:
        with pytest.raises(_native.NativeError) as e:
            ctx.map_split(0, b"text is not records")
        assert e.value.code == _native.MR_ERR_INVALID
    finally:
        ctx.close()
    text = _native.Context(10, 1 << 20, max_keys=1 << 16)
    try:
        with pytest.raises(_native.NativeError):
            text.map_records(0, np.zeros(4, np.uint32), np.zeros(4, np.uint32))
    finally:
        text.close()
    with pytest.raises(_native.NativeError) as e:       # records with the string partitioner: not built
        _native.Context(10, 1 << 20, map_kind=_native.MR_MAP_RECORDS_U32)
    assert e.value.code == _native.MR_ERR_UNSUPPORTED


@pytest.mark.parametrize("kind,op", [(_native.MR_REDUCE_MIN_U64, 1), (_native.MR_REDUCE_MAX_U64, 2),
                                     (_native.MR_REDUCE_COUNT, 3), (_native.MR_REDUCE_SUM_U64, 0)])
def test_other_reducers(kind, op):
    # f-3: min / max / count besides sum, same key order and partitioning
    n = 400_000
    keys, vals = skewed(n, 11, distinct=30_000)
    vals = (vals * np.uint32(4_000_003)).astype(np.uint32)          # spread over the whole uint32 range
    vals[::1000] = 0
    vals[1::1000] = 0xFFFFFFFF
    want = O.groupby_u32([(keys, vals)], 10, op)
    ctx = ctx_for(10, n, reduce_kind=kind)
    try:
        for i, (a, b) in enumerate(zip(np.array_split(keys, 3), np.array_split(vals, 3))):
            ctx.map_records(i, a, b)
        got = ctx.rows()
        assert_rows_equal(got, want, f"reduce kind {kind}")
        ctx.job_begin(keep_input=True)
        assert_rows_equal(ctx.rows(), want, f"reduce kind {kind}, keep_input")
    finally:
        ctx.close()
    if op == 3:
        assert int(got.count.sum()) == n


def test_min_and_max_need_values_to_fold():
    # on a text context they are fed (key, value) pairs (tests/test_gpu_run_files.py); a tokenised split
    # only emits ones, and bigram contexts have no pair input at all
    ctx = _native.Context(10, 1 << 20, reduce_kind=_native.MR_REDUCE_MAX_U64)
    try:
        with pytest.raises(_native.NativeError) as e:
            ctx.map_split(0, b"a b")
        assert e.value.code == _native.MR_ERR_UNSUPPORTED
    finally:
        ctx.close()
    with pytest.raises(_native.NativeError) as e:
        _native.Context(10, 1 << 20, reduce_kind=_native.MR_REDUCE_MAX_U64, map_kind=_native.MR_MAP_BIGRAM_WS)
    assert e.value.code == _native.MR_ERR_UNSUPPORTED
