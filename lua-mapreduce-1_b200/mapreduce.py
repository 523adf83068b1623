"""Host-side mirror of lua-mapreduce's public API for the accelerated path.

The reference's host language is Lua 5.2 (absent from this image), so the host side
above the C ABI is written in Python with the reference's names and call shapes:

    s = mapreduce.server.new(connection_string, dbname)        # init.lua:25-33, server.lua:614
    s.configure{taskfn=..., mapfn=..., partitionfn=..., reducefn=..., finalfn=...,
                combinerfn=..., init_args=..., storage=..., result_ns=...}   # server.lua:417-460
    s.loop()                                                    # server.lua:464-609
    w = mapreduce.worker.new(connection_string, dbname)         # worker.lua:154
    w.configure{max_iter=..., max_sleep=..., max_tasks=...}     # worker.lua:142-148
    w.execute()                                                 # worker.lua:112-138

Task modules keep the reference's fields (examples/WordCount/init.lua:51-64): init,
taskfn, mapfn, partitionfn, reducefn, combinerfn, finalfn and the associative /
commutative / idempotent flags.  A module that is GPU-eligible adds a `native`
descriptor; its Lua/Python functions stay the semantic definition and are NOT
interpreted on the data path.  A job whose modules carry no descriptor is refused
loudly: there is no CPU fallback.

MongoDB is the reference's control plane and is out of scope: the job
board below is an in-process stand-in with the same job/task states (utils.lua:33-46)
so that server and worker keep their roles; connection_string only names the board.
"""
import importlib
import threading
import time

from . import _native

# job states (utils.lua:33-40) and task states (utils.lua:41-46)
WAITING, RUNNING, BROKEN, FINISHED, WRITTEN, FAILED = range(6)
TASK_WAIT, TASK_MAP, TASK_REDUCE, TASK_FINISHED = "WAIT", "MAP", "REDUCE", "FINISHED"
MAX_JOB_RETRIES = 3      # utils.lua:47
MAX_WORKER_RETRIES = 3   # utils.lua:48
DEFAULT_SLEEP = 0.01

NATIVE_KINDS = {
    "map": {"tokenize_ws": _native.MR_MAP_TOKENIZE_WS, "records_u32": _native.MR_MAP_RECORDS_U32,
            "bigram_ws": _native.MR_MAP_BIGRAM_WS},
    "reduce": {"sum_u64": _native.MR_REDUCE_SUM_U64, "min_u64": _native.MR_REDUCE_MIN_U64,
               "max_u64": _native.MR_REDUCE_MAX_U64, "count": _native.MR_REDUCE_COUNT},
    "partition": {"fnv1_lua52_double": _native.MR_PARTITION_FNV1_LUA52, "mod_u32": _native.MR_PARTITION_MOD_U32},
}


class _Board:
    """What the `task`, `map_jobs`, `red_jobs` and `errors` collections hold (Appendix B)."""

    def __init__(self):
        self.lock = threading.Condition()
        self.task = {"status": TASK_WAIT, "iteration": 0, "params": None}
        self.map_jobs, self.red_jobs, self.errors = [], [], []
        self.results = {}   # result name -> cursor factory
        self.finished = False


_boards = {}
_boards_lock = threading.Lock()


def _board(connection_string, dbname):
    with _boards_lock:
        return _boards.setdefault((connection_string, dbname), _Board())


def _load_module(spec):
    """A task module: dotted name (like Lua's require), module, dict or object."""
    m = importlib.import_module(spec) if isinstance(spec, str) else spec
    if isinstance(m, dict):
        return type("task_module", (), m)
    return m


def _field(m, name):
    return getattr(m, name, None)


def _shared_dir(storage):
    """utils.get_storage_from (utils.lua:273-285): "shared:/path" names a directory every worker sees."""
    if isinstance(storage, str) and storage.startswith("shared:"):
        return storage[len("shared:"):]
    return None


def _make_job(key, value):
    # utils.lua:87-98
    return {"_id": key, "value": value, "worker": None, "status": WAITING, "repetitions": 0,
            "creation_time": time.time()}


class server:
    """mapreduce/server.lua, orchestration only."""

    @classmethod
    def new(cls, connection_string, dbname, auth_table=None):
        return cls(connection_string, dbname)

    def __init__(self, connection_string, dbname):
        self.board = _board(connection_string, dbname)
        self.configured = False
        self.finished = False
        self.stats = {}

    def configure(self, params):
        # server.lua:425-443: every role must name a module that has the function and init
        self.mods = {}
        for role in ("taskfn", "mapfn", "partitionfn", "reducefn", "finalfn"):
            assert params.get(role) is not None, f"Field {role} is mandatory"
        for role in ("taskfn", "mapfn", "partitionfn", "reducefn", "finalfn", "combinerfn"):
            spec = params.get(role)
            if spec is None:
                continue
            m = _load_module(spec)
            assert callable(_field(m, role)), f"Module for {role} must have a {role} function"
            assert callable(_field(m, "init")), f"Module for {role} must have an init function"
            self.mods[role] = m
        self.init_args = params.get("init_args")
        self.result_ns = params.get("result_ns", "result")
        self.storage = params.get("storage", "hbm")
        for role in ("taskfn", "finalfn"):   # server.lua:452-456
            self.mods[role].init(self.init_args)
        self.configured = True
        return self

    def loop(self):
        assert self.configured, "configure the server first"
        b = self.board
        it = 0
        while True:
            it += 1
            t0 = time.time()
            with b.lock:
                b.map_jobs, b.red_jobs, b.results = [], [], {}
                b.task = {"status": TASK_WAIT, "iteration": it,
                          "params": {"mods": self.mods, "init_args": self.init_args,
                                     "result_ns": self.result_ns, "storage": self.storage}}
                # server_prepare_map, server.lua:249-276
                keys = set()

                def emit(key, value):
                    assert key not in keys, "taskfn must emit unique keys"
                    keys.add(key)
                    b.map_jobs.append(_make_job(key, value))

                self.mods["taskfn"].taskfn(emit)
                b.task["status"] = TASK_MAP
                b.lock.notify_all()
            self._wait(b.map_jobs)
            t_map = time.time()
            with b.lock:
                # server_prepare_reduce, server.lua:279-329: one job per partition that has keys
                b.task["status"] = TASK_REDUCE
                b.lock.notify_all()
            self._wait_reduce_ready()
            self._wait(b.red_jobs)
            t_red = time.time()
            # server_final, server.lua:346-411: result files in name order = partition order
            names = sorted(b.results)

            def pairs_iterator():
                for name in names:
                    for key, count in b.results[name]():
                        yield key, [count]

            rc = self.mods["finalfn"].finalfn(pairs_iterator())
            self.stats = {"iteration": it, "map_real_time": t_map - t0, "reduce_real_time": t_red - t_map,
                          "server_time": time.time() - t0,
                          "map_jobs": len(b.map_jobs), "red_jobs": len(b.red_jobs),
                          "failed_map_jobs": sum(j["status"] == FAILED for j in b.map_jobs),
                          "failed_red_jobs": sum(j["status"] == FAILED for j in b.red_jobs)}
            if rc == "loop":      # server.lua:386-402
                continue
            with b.lock:
                b.task["status"] = TASK_FINISHED
                b.finished = True
                if rc is True:
                    b.results = {}
                b.lock.notify_all()
            self.finished = True
            return self.stats

    def _wait(self, jobs):
        # make_task_coroutine_wrap, server.lua:186-234: BROKEN x3 -> FAILED; WRITTEN|FAILED are done
        b = self.board
        with b.lock:
            while True:
                for j in jobs:
                    if j["status"] == BROKEN and j["repetitions"] >= MAX_JOB_RETRIES:
                        j["status"] = FAILED
                if all(j["status"] in (WRITTEN, FAILED) for j in jobs):
                    break
                b.lock.wait(DEFAULT_SLEEP)
            if b.errors:
                errs, b.errors = b.errors, []
                if any(j["status"] == FAILED for j in jobs):
                    raise RuntimeError("job failed: " + "; ".join(e["msg"] for e in errs))

    def _wait_reduce_ready(self):
        b = self.board
        with b.lock:
            while b.task.get("reduce_ready") is not True:
                b.lock.wait(DEFAULT_SLEEP)


class worker:
    """mapreduce/worker.lua + the data-plane half of mapreduce/job.lua, on one GPU."""

    @classmethod
    def new(cls, connection_string, dbname, auth_table=None):
        return cls(connection_string, dbname)

    def __init__(self, connection_string, dbname):
        self.board = _board(connection_string, dbname)
        self.max_iter, self.max_sleep, self.max_tasks = 20, 20, 1   # worker.lua:161-163
        self.device = 0
        self.arena_bytes = 1 << 28
        self.max_keys = 0
        self.name = f"worker-{id(self):x}"
        self.ctx = None

    def configure(self, t):
        for k, v in t.items():
            assert k in ("max_iter", "max_sleep", "max_tasks", "device", "arena_bytes", "max_keys"), \
                f"Unknown parameter: {k}"
            setattr(self, k, v)
        return self

    def execute(self):
        # worker.lua:112-138: retry the whole loop up to MAX_WORKER_RETRIES times
        failures = 0
        while True:
            try:
                return self._execute()
            except Exception as e:
                failures += 1
                with self.board.lock:
                    self.board.errors.append({"worker": self.name, "msg": repr(e)})
                    self.board.lock.notify_all()
                if failures >= MAX_WORKER_RETRIES:
                    raise

    # -- native dispatch -------------------------------------------------------
    def _native_config(self, mods):
        desc = None
        for role in ("mapfn", "partitionfn", "reducefn"):
            d = _field(mods[role], "native")
            if d is None:
                raise NotImplementedError(
                    f"task module for {role} has no `native` descriptor: the GPU data plane does not "
                    "interpret mapfn/partitionfn/reducefn and there is no CPU fallback")
            desc = dict(desc or {}, **d)
        for kind, table in NATIVE_KINDS.items():
            if desc.get(kind) not in table:
                raise NotImplementedError(f"native {kind} kind {desc.get(kind)!r} is not built")
        return desc

    def _execute(self):
        b = self.board
        ntasks, it, sleep = 0, 0, DEFAULT_SLEEP
        while it < self.max_iter and ntasks < self.max_tasks:
            with b.lock:
                task = b.task
                if b.finished:
                    return ntasks
                # nothing to take: no task yet, or every job of this iteration is already done by
                # this worker and the server is still in its final phase (a "loop" iteration
                # replaces b.task) -- the reference's take_next_job returns nothing then
                if task["status"] in (TASK_WAIT, TASK_FINISHED) or task is getattr(self, "_done_task", None):
                    it += 1
                    b.lock.wait(min(sleep, self.max_sleep))
                    sleep *= 1.5    # worker.lua:100-101 back-off
                    continue
            self._run_task()
            ntasks += 1
            it, sleep = 0, DEFAULT_SLEEP
        return ntasks

    def _take(self, jobs):
        # task.lua:258-343 take_next_job: first WAITING|BROKEN job becomes RUNNING for this worker
        b = self.board
        with b.lock:
            for j in jobs:
                if j["status"] in (WAITING, BROKEN) and j["repetitions"] < MAX_JOB_RETRIES:
                    j["status"], j["worker"], j["started_time"] = RUNNING, self.name, time.time()
                    return j
        return None

    def _mark(self, job, status, **fields):
        with self.board.lock:
            job["status"] = status
            job.update(fields)
            self.board.lock.notify_all()

    def _broken(self, job, exc):
        # job.lua:322-342 mark_as_broken + worker.lua:129 error text
        with self.board.lock:
            job["status"] = BROKEN
            job["repetitions"] += 1
            job["broken_time"] = time.time()
            self.board.errors.append({"worker": self.name, "msg": repr(exc)})
            self.board.lock.notify_all()

    def _run_task(self):
        b = self.board
        # one iteration = one task document with its own job lists; a "loop" iteration of the
        # server swaps all of them, so this run keeps the ones it started with
        with b.lock:
            task, map_jobs, red_jobs, results = b.task, b.map_jobs, b.red_jobs, b.results
            # ONE worker per board runs the data plane of a task.  The reference's N workers meet
            # through run files in shared storage (job.lua:208-221, 253-260); a worker here keeps
            # its combined records in its own GPU context, and nothing moves them between the
            # contexts of two workers of one board -- each would reduce and publish only what it
            # mapped itself.  So a second worker is refused, loudly, before it touches a job.
            # (N GPUs: N ranks of ONE worker group with a shuffle, parallel.py / mr_shuffle_*.)
            shared_dir = _shared_dir(task["params"].get("storage"))
            owner = self.name if shared_dir else task.setdefault("data_worker", self.name)
            if owner != self.name:
                raise RuntimeError(
                    f"{self.name}: worker {owner} already runs this task's data plane; a board takes one "
                    "worker (several workers would each reduce only their own splits). Use one worker "
                    "per board, or ranks of one shuffle group for several GPUs")
        params = task["params"]
        mods = params["mods"]
        desc = self._native_config(mods)
        R = int(desc["num_reducers"])
        kinds = (NATIVE_KINDS["map"][desc["map"]], NATIVE_KINDS["partition"][desc["partition"]],
                 NATIVE_KINDS["reduce"][desc["reduce"]])
        if self.ctx is None or self.ctx.num_reducers != R or self.ctx_kinds != kinds:
            if self.ctx is not None:
                self.ctx.close()
            self.ctx = _native.Context(R, self.arena_bytes, self.max_keys, device=self.device,
                                       map_kind=kinds[0], partition_kind=kinds[1], reduce_kind=kinds[2])
            self.ctx_kinds = kinds
        ctx = self.ctx
        records = kinds[0] == _native.MR_MAP_RECORDS_U32
        if shared_dir:
            if records or kinds[0] != _native.MR_MAP_TOKENIZE_WS:
                raise NotImplementedError("storage 'shared:' carries string-key runs (word count); number keys and bigrams keep storage 'hbm'")
            return self._run_task_shared(task, map_jobs, red_jobs, results, ctx, R, shared_dir)
        # Map affinity across iterations (task.lua:279-293: in iteration > 1 a worker prefers the map
        # jobs it ran before): when the waiting map jobs are exactly the ones whose splits this
        # worker still holds in HBM, the splits are not staged again (mr_job_begin(keep_input)).
        def signature(j):
            v = j["value"]
            return (j["_id"], v if isinstance(v, (str, bytes)) else id(v))

        with b.lock:
            wanted = [signature(j) for j in map_jobs]
        resident = (task["iteration"] > 1 and getattr(self, "_resident", None) == (id(ctx), wanted)
                    and len(wanted) > 0)
        ctx.job_begin(keep_input=resident)
        self.kept_input = resident
        # ---- map jobs (job.lua:166-227): stage every split this worker claims
        mine = []
        while True:
            job = self._take(map_jobs)
            if job is None:
                with b.lock:
                    if all(j["status"] in (FINISHED, WRITTEN, FAILED) for j in map_jobs):
                        break
                    b.lock.wait(DEFAULT_SLEEP)
                continue
            t0, c0 = time.time(), time.process_time()
            try:
                value = job["value"]
                if resident:    # the split is already in HBM from the previous iteration
                    pass
                elif records:   # the map job's value is a (keys, values) pair of uint32 arrays
                    ctx.map_records(len(mine), value[0], value[1])
                elif isinstance(value, (bytes, bytearray, memoryview)) or hasattr(value, "ctypes"):
                    ctx.map_split(len(mine), value)
                else:   # mapfn.lua:4: the value names a file; the library reads it (mr_map_split_path)
                    ctx.map_split_path(len(mine), value)
                self._mark(job, FINISHED, finished_time=time.time())
                mine.append((job, t0, c0))
            except Exception as e:
                self._broken(job, e)
        try:
            ctx.map_finish()
            ctx.reduce_all()
            info = ctx.job_sync()
        except Exception as e:
            for job, _, _ in mine:
                self._broken(job, e)
            raise
        for job, t0, c0 in mine:   # job.lua:134-152 mark_as_written with cpu_time / real_time
            self._mark(job, WRITTEN, written_time=time.time(), cpu_time=time.process_time() - c0,
                       real_time=time.time() - t0)
        self.last_info = info
        self._resident = (id(ctx), wanted) if len(mine) == len(wanted) else None
        # ---- wait for the reduce phase, publish one reduce job per non-empty partition
        with b.lock:
            while task["status"] == TASK_MAP:
                b.lock.wait(DEFAULT_SLEEP)
            if not task.get("reduce_ready"):
                rows_per_part = self._partition_sizes(ctx)
                width = len(str(max(R - 1, 0)))
                for p in range(R):   # server.lua:297-324: empty partitions get no job, no file
                    if rows_per_part[p]:
                        name = f"{params['result_ns']}.P{p:0{width}d}"
                        red_jobs.append(_make_job(p, {"partition": p, "result": name}))
                task["reduce_ready"] = True
                b.lock.notify_all()
        # ---- reduce jobs (job.lua:232-295): rows are already merged and sorted in HBM
        while True:
            job = self._take(red_jobs)
            if job is None:
                break
            t0, c0 = time.time(), time.process_time()
            try:
                p = job["value"]["partition"]
                rows = list(ctx.partition_cursor(p))
                if records:     # number keys travel as 4 big-endian bytes
                    rows = [(int.from_bytes(k, "big"), c) for k, c in rows]
                with b.lock:
                    results[job["value"]["result"]] = (lambda r=rows: iter(r))
                self._mark(job, WRITTEN, written_time=time.time(), cpu_time=time.process_time() - c0,
                           real_time=time.time() - t0)
            except Exception as e:
                self._broken(job, e)
        with b.lock:
            while not all(j["status"] in (WRITTEN, FAILED) for j in red_jobs):
                b.lock.wait(DEFAULT_SLEEP)
        self._done_task = task

    def _run_task_shared(self, task, map_jobs, red_jobs, results, ctx, R, path):
        """storage = "shared:<dir>": the reference's own arrangement (fs.lua 'shared', job.lua:208-221 and
        253-265).  Any number of workers share the board: a worker maps the splits it claims and leaves
        one run `map_results.P<p>.M<worker>` per partition in <dir>; reduce jobs, one per partition that
        has runs (server.lua:294-324), go to whichever worker takes them: it reads all runs of its
        partition into its context (mr_map_pairs), reduces, publishes the result and removes the runs."""
        import os
        from . import wire
        b = self.board
        params = task["params"]
        ns = "map_results.it%d" % task["iteration"]
        ctx.job_begin()
        mine = []
        while True:
            job = self._take(map_jobs)
            if job is None:
                with b.lock:
                    if all(j["status"] in (FINISHED, WRITTEN, FAILED) for j in map_jobs):
                        break
                    b.lock.wait(DEFAULT_SLEEP)
                continue
            t0, c0 = time.time(), time.process_time()
            try:
                value = job["value"]
                if isinstance(value, (bytes, bytearray, memoryview)) or hasattr(value, "ctypes"):
                    ctx.map_split(len(mine), value)
                else:
                    ctx.map_split_path(len(mine), value)
                self._mark(job, FINISHED, finished_time=time.time())
                mine.append((job, t0, c0))
            except Exception as e:
                self._broken(job, e)
        self.kept_input = False
        if mine:
            try:
                ctx.map_finish()
                ctx.reduce_all()
                self.last_info = ctx.job_sync()
                wire.write_map_results(ctx.rows(), path, mapper_key=self.name, map_results_ns=ns)
            except Exception as e:
                for job, _, _ in mine:
                    self._broken(job, e)
                raise
            for job, t0, c0 in mine:
                self._mark(job, WRITTEN, written_time=time.time(), cpu_time=time.process_time() - c0,
                           real_time=time.time() - t0)
        # ---- reduce phase: the first worker to see it lists the partitions that have runs
        with b.lock:
            while task["status"] == TASK_MAP:
                b.lock.wait(DEFAULT_SLEEP)
            if not task.get("reduce_ready"):
                parts = wire.partitions_with_runs(path, ns)
                width = len(str(max(parts))) if parts else 1
                for p in parts:
                    name = f"{params['result_ns']}.P{p:0{width}d}"
                    red_jobs.append(_make_job(p, {"partition": p, "result": name}))
                task["reduce_ready"] = True
                b.lock.notify_all()
        while True:
            job = self._take(red_jobs)
            if job is None:
                break
            t0, c0 = time.time(), time.process_time()
            try:
                p = job["value"]["partition"]
                names = wire.list_map_results(path, p, ns)
                ctx.job_begin()
                wire.ingest_runs(ctx, path, p, ns)
                ctx.reduce_all()
                ctx.job_sync()
                rows = list(ctx.partition_cursor(p))
                with b.lock:
                    results[job["value"]["result"]] = (lambda r=rows: iter(r))
                for n in names:                       # job.lua:293: the consumed runs are deleted
                    os.remove(os.path.join(path, n))
                self._mark(job, WRITTEN, written_time=time.time(), cpu_time=time.process_time() - c0,
                           real_time=time.time() - t0)
            except Exception as e:
                self._broken(job, e)
        with b.lock:
            while not all(j["status"] in (WRITTEN, FAILED) for j in red_jobs):
                b.lock.wait(DEFAULT_SLEEP)
        self._done_task = task

    @staticmethod
    def _partition_sizes(ctx):
        import numpy as np
        rows = ctx.rows()
        return np.diff(rows.part_begin.astype(np.int64)).tolist()


def utest():
    """mapreduce/test.lua runs every module's utest(); the pure-host pieces live here."""
    assert _make_job(1, "x")["status"] == WAITING
    b = _board("utest", "utest")
    assert b is _board("utest", "utest")
