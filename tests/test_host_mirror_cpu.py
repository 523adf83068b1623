"""The server/worker mirror's control flow on CPU: a stand-in for _native.Context (the oracle plays
the device; this is test code) lets the job board, the BROKEN -> retry -> FAILED rules, the
"loop" iterations with map affinity and the worker's bookkeeping run without a GPU.  The data plane
itself is covered by the -m gpu tests."""
import os
import threading
import time
import types

import numpy as np
import pytest

from lua_mapreduce_1_b200 import mapreduce
from lua_mapreduce_1_b200.examples import WordCount
from oracle import oracle as O


class StandInContext:
    """What mapreduce.worker calls on a context, answered by the oracle."""
    created = 0
    fail_next_splits = 0            # class-wide: that many map_split calls raise (a BROKEN job each)
    default_delay = 0.0             # class-wide: seconds every reduce_all takes

    def __init__(self, num_reducers, arena_bytes, max_keys=0, **kw):
        type(self).created += 1
        self.num_reducers = num_reducers
        self.splits, self.rows_, self.h2d = [], None, 0
        self.delay = type(self).default_delay

    def job_begin(self, keep_input=False):
        if not keep_input:
            self.splits = []
        self.h2d = 0
        self.rows_ = None
        self.pairs = {}

    def map_pairs(self, split_id, keys, values):
        for k, v in zip(keys, values):
            self.pairs[k] = self.pairs.get(k, 0) + int(v)

    def _stage(self, data):
        if type(self).fail_next_splits > 0:
            type(self).fail_next_splits -= 1
            raise RuntimeError("staging failed (injected)")
        self.splits.append(bytes(data))
        self.h2d += len(data)

    def map_split(self, split_id, data):
        self._stage(data.tobytes() if hasattr(data, "tobytes") else data)

    def map_split_path(self, split_id, path):
        with open(path, "rb") as f:
            self._stage(f.read())

    def map_finish(self):
        pass

    def reduce_all(self):
        time.sleep(self.delay)
        self.rows_ = O.run_naive(self.splits, self.num_reducers)
        if getattr(self, "pairs", None):        # ingested runs merge with the tokenised splits (mr_map_pairs)
            merged = dict(zip(self.rows_.keys(), (int(c) for c in self.rows_.count)))
            for k, v in self.pairs.items():
                merged[k] = merged.get(k, 0) + v
            order = sorted(merged, key=lambda k: (O.partition(k, self.num_reducers), k))
            off = np.concatenate([[0], np.cumsum([len(k) for k in order])]).astype(np.uint64)
            self.rows_ = O.Rows([O.partition(k, self.num_reducers) for k in order], [merged[k] for k in order], off,
                                np.frombuffer(b"".join(order), dtype=np.uint8), self.rows_.tokens + len(self.pairs))

    def job_sync(self):
        return types.SimpleNamespace(n_tokens=self.rows_.tokens, n_keys=self.rows_.n)

    def rows(self):
        pb = np.searchsorted(self.rows_.part, np.arange(self.num_reducers + 1))
        r = self.rows_
        return types.SimpleNamespace(part_begin=pb.astype(np.uint64), part=r.part, count=r.count, key_off=r.key_off,
                                     key_bytes=r.key_bytes)

    def partition_cursor(self, p):
        for part, key, count in zip(self.rows_.part, self.rows_.keys(), self.rows_.count):
            if int(part) == p:
                yield key, int(count)

    def stats(self):
        return types.SimpleNamespace(h2d_bytes=self.h2d)

    def close(self):
        pass


@pytest.fixture
def stand_in(monkeypatch):
    StandInContext.created = 0
    StandInContext.fail_next_splits = 0
    StandInContext.default_delay = 0.0
    monkeypatch.setattr(mapreduce._native, "Context", StandInContext)
    return StandInContext


def write_splits(tmp_path, texts):
    paths = []
    for i, t in enumerate(texts):
        p = tmp_path / f"s{i}.txt"
        p.write_bytes(t)
        paths.append(str(p))
    return paths


def run(db, paths, final, worker_cfg=None, R=15):
    s = mapreduce.server.new("localhost", db)
    s.configure({"taskfn": WordCount, "mapfn": WordCount, "partitionfn": WordCount, "reducefn": WordCount,
                 "combinerfn": WordCount, "finalfn": final, "init_args": {"files": paths, "num_reducers": R}})
    w = mapreduce.worker.new("localhost", db).configure(dict({"max_iter": 2000, "max_tasks": 1}, **(worker_cfg or {})))
    t = threading.Thread(target=w.execute, daemon=True)
    t.start()
    stats = s.loop()
    t.join(30)
    assert not t.is_alive()
    return s, w, stats


def final_collect(runs, loops=1):
    class Final:
        init = staticmethod(lambda arg=None: None)

        @staticmethod
        def finalfn(pairs_iterator):
            runs.append([(k, v[0]) for k, v in pairs_iterator])
            return "loop" if len(runs) < loops else True
    return Final


def test_one_task_reaches_finalfn_in_the_reference_order(tmp_path, stand_in):
    texts = [b"b a c a\n", b"", b"a  b\tz\n", b"q"]
    want = O.run_naive(texts, 15)
    runs = []
    s, w, stats = run("cpu-one", write_splits(tmp_path, texts), final_collect(runs))
    assert runs == [list(zip(want.keys(), (int(c) for c in want.count)))]       # partition asc, key asc
    assert stats["map_jobs"] == 4 and stats["failed_map_jobs"] == 0
    assert stats["red_jobs"] == len(set(int(p) for p in want.part))              # no job for an empty partition
    assert stand_in.created == 1


@pytest.mark.parametrize("delay", [0.0, 0.02])
def test_loop_iterations_never_rerun_a_finished_iteration(tmp_path, stand_in, delay):
    # the worker is done with an iteration while the server is still in finalfn: it must wait for the
    # next task document instead of running the old one again (that used to deadlock the pair)
    texts = [b"x y x\n", b"y z\n", b"x\n"]
    want = O.run_naive(texts, 15)
    expect = list(zip(want.keys(), (int(c) for c in want.count)))
    runs = []
    final = final_collect(runs, loops=4)
    slow_final = final.finalfn

    def finalfn(it):
        rc = slow_final(it)
        time.sleep(delay)                  # the window in which the worker sees TASK_REDUCE with nothing left to do
        return rc
    final.finalfn = staticmethod(finalfn)
    s, w, stats = run("cpu-loop-%s" % delay, write_splits(tmp_path, texts), final, {"max_tasks": 4})
    assert stats["iteration"] == 4 and runs == [expect] * 4
    assert w.kept_input is True and int(w.ctx.stats().h2d_bytes) == 0            # later iterations staged nothing
    assert stand_in.created == 1


def test_a_broken_map_job_is_retried_and_a_hopeless_one_fails_the_task(tmp_path, stand_in):
    texts = [b"a b\n", b"c\n"]
    want = O.run_naive(texts, 15)
    stand_in.fail_next_splits = 2          # two failures, third try succeeds (MAX_JOB_RETRIES = 3)
    runs = []
    s, w, stats = run("cpu-retry", write_splits(tmp_path, texts), final_collect(runs))
    assert runs == [list(zip(want.keys(), (int(c) for c in want.count)))]
    assert stats["failed_map_jobs"] == 0
    stand_in.fail_next_splits = 10**6      # never succeeds: BROKEN x3 -> FAILED, the server reports it
    s2 = mapreduce.server.new("localhost", "cpu-fail")
    s2.configure({"taskfn": WordCount, "mapfn": WordCount, "partitionfn": WordCount, "reducefn": WordCount,
                  "combinerfn": WordCount, "finalfn": final_collect([]), "init_args": {"files": write_splits(tmp_path, texts),
                                                                                       "num_reducers": 15}})
    w2 = mapreduce.worker.new("localhost", "cpu-fail").configure({"max_iter": 2000, "max_tasks": 1})
    t = threading.Thread(target=w2.execute, daemon=True)
    t.start()
    with pytest.raises(RuntimeError, match="job failed"):
        s2.loop()


def test_a_second_worker_on_the_board_is_refused_and_the_result_is_whole(tmp_path, stand_in):
    # the reference runs N workers per database (worker.lua, task.lua:258-343); here a worker keeps its
    # combined records in its own GPU context, so two workers would each publish partial partitions.
    # The second one must fail loudly before it claims a job, and the first one's result must be complete.
    texts = [b"a b a\n", b"b c\n", b"c d a\n", b"e\n"] * 6
    want = O.run_naive(texts, 15)
    runs = []
    stand_in.default_delay = 0.3           # the first worker is still inside the task when the second arrives
    s = mapreduce.server.new("localhost", "cpu-two-workers")
    s.configure({"taskfn": WordCount, "mapfn": WordCount, "partitionfn": WordCount, "reducefn": WordCount,
                 "combinerfn": WordCount, "finalfn": final_collect(runs),
                 "init_args": {"files": write_splits(tmp_path, texts), "num_reducers": 15}})
    workers = [mapreduce.worker.new("localhost", "cpu-two-workers").configure({"max_iter": 2000, "max_tasks": 1})
               for _ in range(2)]
    errors = []

    def go(w):
        try:
            w.execute()
        except Exception as e:
            errors.append(e)
    threads = [threading.Thread(target=go, args=(w,), daemon=True) for w in workers]
    for t in threads:
        t.start()
    s.loop()
    for t in threads:
        t.join(30)
        assert not t.is_alive()
    assert runs == [list(zip(want.keys(), (int(c) for c in want.count)))]       # whole, not one worker's share
    assert len(errors) == 1 and "already runs this task's data plane" in str(errors[0])
    assert stand_in.created == 1                                                 # the refused worker never made a context


def test_many_workers_share_a_board_through_run_files(tmp_path, stand_in):
    # storage "shared:<dir>": the reference's deployment -- N workers per board, mappers and reducers meet
    # through map_results.P<p>.M<worker> files (job.lua:208-221, 253-265).  Whoever claims what, the result
    # is whole, every reduce job read every run of its partition, and the runs are gone afterwards.
    texts = [b"a b a\n", b"b c\n", b"c d a\n", b"e \"q\" back\\slash\n", b"", b"a " * 50] * 5
    want = O.run_naive(texts, 15)
    expect = list(zip(want.keys(), (int(c) for c in want.count)))
    store = tmp_path / "runs"
    stand_in.default_delay = 0.01
    for loops in (1, 2):
        runs = []
        db = f"cpu-shared-{loops}"
        s = mapreduce.server.new("localhost", db)
        s.configure({"taskfn": WordCount, "mapfn": WordCount, "partitionfn": WordCount, "reducefn": WordCount,
                     "combinerfn": WordCount, "finalfn": final_collect(runs, loops=loops), "storage": f"shared:{store}",
                     "init_args": {"files": write_splits(tmp_path, texts), "num_reducers": 15}})
        workers = [mapreduce.worker.new("localhost", db).configure({"max_iter": 4000, "max_tasks": loops}) for _ in range(3)]
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
            t.join(30)
            assert not t.is_alive()
        assert not errors, errors
        assert runs == [expect] * loops
        assert stats["red_jobs"] == len(set(int(p) for p in want.part))
        assert [n for n in os.listdir(store) if n.startswith("map_results")] == []      # consumed runs are removed
