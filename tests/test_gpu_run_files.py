"""Run-file interchange: the reference's mappers and reducers meet through text runs
`<path>/map_results.P<part>.M<mapper>` of `return <%q key>,{<values>}` lines (job.lua:208-221 write,
job.lua:253-265 + utils.merge_iterator read).  A GPU mapper must be able to write them and a GPU reducer
to read them -- including runs a reference mapper wrote, with value lists longer than one."""
import os

import numpy as np
import pytest

from helpers import assert_rows_equal, merge_iterator
from lua_mapreduce_1_b200 import _native, corpus, wire
from oracle import oracle as O

pytestmark = pytest.mark.gpu
R = 10


def reference_style_runs(splits, directory, first_mapper=0):
    """What reference map jobs would leave: per split one run per partition, keys ascending; a key seen
    up to 6 times keeps its values as a list of ones (the inline combiner only folds lists longer than
    MAX_MAP_RESULT, job.lua:92-96, and the final one is skipped for... a single value, job.lua:198)."""
    for j, s in enumerate(splits):
        rows = O.run_naive([s], R)
        keys = rows.keys()
        for p in range(R):
            idx = [i for i in range(rows.n) if int(rows.part[i]) == p]
            if not idx:
                continue
            with open(os.path.join(directory, wire.map_result_name("map_results", p, first_mapper + j)), "wb") as f:
                for i in idx:
                    c = int(rows.count[i])
                    f.write(wire.format_line(keys[i], [1] * c if c <= 6 else [c]))


def texts():
    splits, _ = corpus.generate(lines=[700, 900, 1, 0, 500, 650])
    splits = [bytes(s) for s in splits]
    splits.append(corpus.TORTURE)                                   # quotes, backslashes, control bytes, 300-byte keys
    splits.append(b"caf\xc3\xa9 caf\xc3\xa9 na\xc3\xafve \"q\" back\\slash " + b"Z" * 40 + b" " + b"Z" * 40)
    return splits


def test_a_gpu_reducer_reads_runs_written_by_reference_style_mappers(tmp_path):
    splits = texts()
    reference_style_runs(splits, str(tmp_path))
    want = O.run_naive(splits, R)
    assert wire.partitions_with_runs(str(tmp_path)) == sorted(set(int(p) for p in want.part))
    ctx = _native.Context(R, 1 << 23, max_keys=1 << 17)
    try:
        pairs = 0
        for p in wire.partitions_with_runs(str(tmp_path)):           # one reduce job per partition that has runs
            nruns, npairs = wire.ingest_runs(ctx, str(tmp_path), p, batch=777)
            assert nruns >= 1
            pairs += npairs
        got = ctx.rows()
        info = ctx.job_sync()
        assert_rows_equal(got, want, "runs in")
        assert int(info.n_tokens) == pairs
        with pytest.raises(_native.NativeError) as e:                # the runs are not kept as text
            ctx.job_begin(keep_input=True)
        assert e.value.code == _native.MR_ERR_UNSUPPORTED
        # one partition alone, as a reduce job of the reference does it (job.lua:253-265)
        ctx.job_begin()
        wire.ingest_runs(ctx, str(tmp_path), 3)
        got3 = ctx.rows()
        sel = want.part == 3
        assert got3.keys() == [k for k, s in zip(want.keys(), sel) if s] and np.array_equal(got3.count, want.count[sel])
    finally:
        ctx.close()


def test_text_splits_and_ingested_runs_merge_in_one_job(tmp_path):
    splits = texts()
    reference_style_runs(splits[::2], str(tmp_path))                 # every other split was mapped elsewhere
    want = O.run_naive(splits, R)
    ctx = _native.Context(R, 1 << 23, max_keys=1 << 17)
    try:
        ctx.map_split(0, splits[1])
        for p in wire.partitions_with_runs(str(tmp_path))[:4]:
            wire.ingest_runs(ctx, str(tmp_path), p)
        for i, s in enumerate(splits[3::2]):
            ctx.map_split(1 + i, s)
        for p in wire.partitions_with_runs(str(tmp_path))[4:]:
            wire.ingest_runs(ctx, str(tmp_path), p)
        assert_rows_equal(ctx.rows(), want, "text + runs")
        with pytest.raises(_native.NativeError):
            ctx.job_begin()
            ctx.map_pairs(0, [b"ok", b""], [1, 2])                   # an empty key is refused
            ctx.rows()
    finally:
        ctx.close()


def test_runs_written_by_gpu_mappers_feed_the_reference_merge(tmp_path):
    splits = texts()
    want = O.run_naive(splits, R)
    for mapper in range(2):                                          # two GPU mappers, each with its own splits
        ctx = _native.Context(R, 1 << 23, max_keys=1 << 17)
        try:
            for i, s in enumerate(splits[mapper::2]):
                ctx.map_split(i, s)
            names = wire.write_map_results(ctx.rows(), str(tmp_path), mapper_key=f"gpu{mapper}")
            assert all(n.startswith("map_results.P") and n.endswith(f".Mgpu{mapper}") for n in names)
        finally:
            ctx.close()
    # the reduce side of the reference, restated: per partition, k-way merge of its runs, sum reducer
    got = []
    for p in wire.partitions_with_runs(str(tmp_path)):
        runs = [wire.read_run(os.path.join(str(tmp_path), n)) for n in wire.list_map_results(str(tmp_path), p)]
        for key, values in merge_iterator(runs):
            got.append((p, key, sum(values)))
    assert got == [(int(p), k, int(c)) for p, k, c in zip(want.part, want.keys(), want.count)]


def test_min_and_max_over_string_keys_with_values():
    # f-3: reducers other than the sum on string keys.  A tokenised split only ever emits (word, 1), so a
    # min/max context is fed (key, value) pairs -- e.g. runs whose lines carry the values to fold
    import random
    rnd = random.Random(6)
    keys = [bytes(rnd.choice(b"abcdefgh") for _ in range(rnd.randrange(1, 24))) for _ in range(3000)]
    pairs = [(rnd.choice(keys), rnd.randrange(0, 1 << 40)) for _ in range(40000)]
    for kind, fold in ((_native.MR_REDUCE_MIN_U64, min), (_native.MR_REDUCE_MAX_U64, max)):
        want = {}
        for k, v in pairs:
            want[k] = fold(want[k], v) if k in want else v
        ctx = _native.Context(R, 1 << 23, max_keys=1 << 16, reduce_kind=kind)
        try:
            with pytest.raises(_native.NativeError) as e:
                ctx.map_split(0, b"text is refused here")
            assert e.value.code == _native.MR_ERR_UNSUPPORTED
            for b in range(0, len(pairs), 7000):
                chunk = pairs[b:b + 7000]
                ctx.map_pairs(b, [k for k, _ in chunk], [v for _, v in chunk])
            rows = ctx.rows()
        finally:
            ctx.close()
        got = dict(zip(rows.keys(), (int(c) for c in rows.count)))
        assert got == want
        order = [(wire_part(k), k) for k in rows.keys()]
        assert order == sorted(order) and [p for p, _ in order] == [int(p) for p in rows.part]
    # a count over tokenised text is the sum of the ones
    ctx = _native.Context(R, 1 << 20, max_keys=1 << 12, reduce_kind=_native.MR_REDUCE_COUNT)
    try:
        ctx.map_split(0, b"a b a c a b")
        rows = ctx.rows()
        assert dict(zip(rows.keys(), (int(c) for c in rows.count))) == {b"a": 3, b"b": 2, b"c": 1}
    finally:
        ctx.close()


def wire_part(key):
    return O.partition(key, R)
