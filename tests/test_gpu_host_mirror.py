"""The reference-shaped host API (server.configure/loop, worker.execute, task modules) on a GPU,
and the reference's result wire format written from GPU rows."""
import threading

import numpy as np
import pytest

from lua_mapreduce_1_b200 import _native, corpus, mapreduce, wire
from lua_mapreduce_1_b200.examples import WordCount
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def test_wordcount_through_server_and_worker(tmp_path, capsys):
    splits, _ = corpus.generate(lines=[700, 650, 1, 900])
    splits[2] = np.frombuffer(corpus.TORTURE, dtype=np.uint8)
    paths = []
    for i, s in enumerate(splits):
        p = tmp_path / f"split{i}.txt"
        p.write_bytes(s.tobytes())
        paths.append(str(p))
    want = O.run_staged(splits, 15, 2)

    seen = []

    class Final:                              # finalfn.lua:3-8, collecting instead of printing
        init = staticmethod(lambda arg=None: None)

        @staticmethod
        def finalfn(pairs_iterator):
            for key, values in pairs_iterator:
                seen.append((key, values[0]))
            return True

    s = mapreduce.server.new("localhost", "wordcount-gpu")
    s.configure({"taskfn": WordCount, "mapfn": WordCount, "partitionfn": WordCount, "reducefn": WordCount,
                 "combinerfn": WordCount, "finalfn": Final, "init_args": {"files": paths, "num_reducers": 15}})
    w = mapreduce.worker.new("localhost", "wordcount-gpu").configure({"max_iter": 2000, "arena_bytes": 1 << 24,
                                                                       "max_keys": 1 << 18})
    err = []
    t = threading.Thread(target=lambda: err.append(None) if w.execute() is not None else err.append("x"))
    t.start()
    stats = s.loop()
    t.join(60)
    assert not t.is_alive()
    # server_final order: partitions ascending, keys ascending (server.lua:365, utils.lua:214)
    assert seen == list(zip(want.keys(), (int(c) for c in want.count)))
    assert stats["map_jobs"] == 4 and stats["failed_map_jobs"] == 0
    assert stats["red_jobs"] == len(set(want.part.tolist()))          # no job for an empty partition
    # the module's own partitionfn (the semantic definition) agrees with the device partition ids
    for p, k in list(zip(want.part, want.keys()))[::97]:
        assert WordCount.partitionfn(k) == int(p)
    w.ctx.close()


def test_results_in_the_reference_wire_format(tmp_path):
    splits, _ = corpus.generate(lines=[500, 400])
    splits.append(np.frombuffer(corpus.TORTURE, dtype=np.uint8))
    ctx = _native.Context(15, 1 << 22, max_keys=1 << 17)
    try:
        for i, s in enumerate(splits):
            ctx.map_split(i, s)
        rows = ctx.rows()
    finally:
        ctx.close()
    names = wire.write_results(rows, str(tmp_path), "result")
    want = O.run_staged(splits, 15, 2)
    parts = sorted(set(want.part.tolist()))
    assert names == [wire.result_name("result", p, parts[-1]) for p in parts]
    got = list(wire.read_pairs(str(tmp_path), names))                  # what server_final hands to finalfn
    assert got == [(k, [int(c)]) for k, c in zip(want.keys(), want.count)]
    first = (tmp_path / names[0]).read_bytes().split(b"\n")[0]
    assert first.startswith(b'return "') and first.endswith(b"}")


def test_group_by_sum_through_server_and_worker():
    from lua_mapreduce_1_b200.examples import GroupBySum
    rng = np.random.default_rng(3)
    batches = [(rng.integers(0, 5000, n).astype(np.uint32), rng.integers(0, 100, n).astype(np.uint32))
               for n in (20000, 1, 35000)]
    want = O.groupby_u32(batches, 7)
    seen = []

    class Final:
        init = staticmethod(lambda arg=None: None)

        @staticmethod
        def finalfn(pairs_iterator):
            seen.extend((k, v[0]) for k, v in pairs_iterator)
            return True

    s = mapreduce.server.new("localhost", "groupby-gpu")
    s.configure({"taskfn": GroupBySum, "mapfn": GroupBySum, "partitionfn": GroupBySum, "reducefn": GroupBySum,
                 "combinerfn": GroupBySum, "finalfn": Final, "init_args": {"batches": batches, "num_reducers": 7}})
    w = mapreduce.worker.new("localhost", "groupby-gpu").configure({"max_iter": 2000, "arena_bytes": 1 << 22,
                                                                     "max_keys": 1 << 16})
    t = threading.Thread(target=w.execute)
    t.start()
    s.loop()
    t.join(60)
    assert not t.is_alive()
    keys = np.frombuffer(want.key_bytes.tobytes(), ">u4").tolist()
    assert seen == list(zip(keys, (int(c) for c in want.count)))
    w.ctx.close()


def test_loop_iterations_keep_the_splits_resident(tmp_path):
    # finalfn returning "loop" (server.lua:386-402) starts another iteration; the worker then finds
    # the same map jobs and maps the splits it already holds in HBM (map affinity, task.lua:279-293)
    splits, _ = corpus.generate(lines=[400, 300])
    paths = []
    for i, s_ in enumerate(splits):
        p = tmp_path / f"s{i}.txt"
        p.write_bytes(s_.tobytes())
        paths.append(str(p))
    want = O.run_staged(splits, 15, 2)
    runs = []

    class Final:
        init = staticmethod(lambda arg=None: None)

        @staticmethod
        def finalfn(pairs_iterator):
            runs.append([(k, v[0]) for k, v in pairs_iterator])
            return "loop" if len(runs) < 3 else True

    s = mapreduce.server.new("localhost", "loop-gpu")
    s.configure({"taskfn": WordCount, "mapfn": WordCount, "partitionfn": WordCount, "reducefn": WordCount,
                 "combinerfn": WordCount, "finalfn": Final, "init_args": {"files": paths, "num_reducers": 15}})
    w = mapreduce.worker.new("localhost", "loop-gpu").configure({"max_iter": 2000, "max_tasks": 3,
                                                                  "arena_bytes": 1 << 22, "max_keys": 1 << 17})
    t = threading.Thread(target=w.execute)
    t.start()
    stats = s.loop()
    t.join(60)
    assert not t.is_alive() and stats["iteration"] == 3
    expect = list(zip(want.keys(), (int(c) for c in want.count)))
    assert runs == [expect, expect, expect]
    assert w.kept_input is True and int(w.ctx.stats().h2d_bytes) == 0     # last iteration copied nothing
    w.ctx.close()


def test_three_gpu_workers_share_a_board_through_run_files(tmp_path):
    # storage "shared:<dir>": N workers per board as in the reference; GPU mappers write the reference's
    # run files, GPU reducers read them back (mr_map_pairs) -- whoever claims what, the result is whole
    splits, _ = corpus.generate(lines=[500, 650, 1, 400, 300, 0, 720, 210])
    splits[2] = np.frombuffer(corpus.TORTURE, dtype=np.uint8)
    paths = []
    for i, s in enumerate(splits):
        p = tmp_path / f"split{i}.txt"
        p.write_bytes(s.tobytes())
        paths.append(str(p))
    want = O.run_staged(splits, 15, 2)
    seen = []

    class Final:
        init = staticmethod(lambda arg=None: None)

        @staticmethod
        def finalfn(pairs_iterator):
            seen.append([(k, v[0]) for k, v in pairs_iterator])
            return True

    store = tmp_path / "store"
    s = mapreduce.server.new("localhost", "wordcount-gpu-shared")
    s.configure({"taskfn": WordCount, "mapfn": WordCount, "partitionfn": WordCount, "reducefn": WordCount,
                 "combinerfn": WordCount, "finalfn": Final, "storage": f"shared:{store}",
                 "init_args": {"files": paths, "num_reducers": 15}})
    workers = [mapreduce.worker.new("localhost", "wordcount-gpu-shared").configure(
        {"max_iter": 4000, "arena_bytes": 1 << 24, "max_keys": 1 << 18}) for _ in range(3)]
    errors = []

    def go(w):
        try:
            w.execute()
        except Exception as e:
            errors.append(e)
    threads = [threading.Thread(target=go, args=(w,), daemon=True) for w in workers]
    for t in threads:
        t.start()
    stats = s.loop()
    for t in threads:
        t.join(60)
        assert not t.is_alive()
    assert not errors, errors
    assert seen == [list(zip(want.keys(), (int(c) for c in want.count)))]
    assert stats["red_jobs"] == len(set(want.part.tolist())) and stats["failed_map_jobs"] == 0
    for w in workers:
        if w.ctx is not None:
            w.ctx.close()
