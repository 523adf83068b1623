"""bench.py's parity digest (the check that makes a multi-GPU bench line mean "bit-exact") on CPU."""
import numpy as np

import bench
from lua_mapreduce_1_b200 import corpus
from oracle import oracle as O


def _slice(rows, idx):
    keys = rows.keys()
    kb = b"".join(keys[i] for i in idx)
    off = np.concatenate([[0], np.cumsum([len(keys[i]) for i in idx])]).astype(np.uint64)
    return O.Rows(rows.part[idx], rows.count[idx], off, np.frombuffer(kb, dtype=np.uint8), 0)


def test_digest_adds_up_over_any_sharding_of_the_rows_and_sees_every_field():
    splits, tok = corpus.generate(lines=[600, 500])
    splits.append(np.frombuffer(corpus.TORTURE, dtype=np.uint8))        # 300-byte keys, NUL bytes
    rows = O.run_naive(splits, 10)
    whole = bench.rows_digest64(rows.part, rows.count, rows.key_off, rows.key_bytes)
    assert whole[1] == rows.n and whole[2] == rows.tokens
    rnd = np.random.RandomState(1)
    owner = rnd.randint(0, 3, size=rows.n)
    parts = [_slice(rows, np.nonzero(owner == r)[0]) for r in range(3)]
    ds = [bench.rows_digest64(p.part, p.count, p.key_off, p.key_bytes) for p in parts]
    assert sum(d[0] for d in ds) % (1 << 64) == whole[0]
    assert sum(d[1] for d in ds) == whole[1] and sum(d[2] for d in ds) == whole[2]
    # a slice whose key_off does not start at zero (a rank's rows inside a bigger buffer)
    sub = np.arange(5, 50)
    a = _slice(rows, sub)
    assert bench.rows_digest64(rows.part[5:50], rows.count[5:50], rows.key_off[5:51], rows.key_bytes) == \
        bench.rows_digest64(a.part, a.count, a.key_off, a.key_bytes)

    def changed(**kw):
        f = {"part": rows.part.copy(), "count": rows.count.copy(), "key_off": rows.key_off.copy(),
             "key_bytes": rows.key_bytes.copy()}
        for k, fn in kw.items():
            fn(f[k])
        return bench.rows_digest64(f["part"], f["count"], f["key_off"], f["key_bytes"])[0]
    assert changed(count=lambda c: c.__setitem__(7, c[7] + 1)) != whole[0]
    assert changed(part=lambda p: p.__setitem__(7, (p[7] + 1) % 10)) != whole[0]
    assert changed(key_bytes=lambda b: b.__setitem__(3, b[3] ^ 1)) != whole[0]
    # moving a byte from one key to its neighbour (same bytes, different boundaries)
    assert changed(key_off=lambda o: o.__setitem__(1, o[1] + 1)) != whole[0]
    assert bench.rows_digest64([], [], [0], []) == (0, 0, 0)


def test_rows_in_order_is_the_reference_order():
    rows = O.run_naive([b"b a aa a\x00 B ab"], 1)
    assert bench.rows_in_order(rows)
    swapped = O.Rows(rows.part, rows.count, rows.key_off, rows.key_bytes[::-1].copy(), 0)
    assert not bench.rows_in_order(swapped)


def test_host_threads_ignores_omp_num_threads(monkeypatch):
    monkeypatch.setenv("OMP_NUM_THREADS", "1")       # what torch.distributed.run injects into every rank
    import os
    assert bench.host_threads() == len(os.sched_getaffinity(0))
