"""Shared helpers for the parity tests (test infrastructure; may use oracle/)."""
import hashlib
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def golden(name):
    with open(os.path.join(HERE, "golden", name)) as f:
        return json.load(f)


def rows_digest(part, count, key_off, key_bytes):
    """sha256 over 'partition<TAB>count<TAB>key<LF>' lines, the format of make_golden.py."""
    h = hashlib.sha256()
    kb = np.asarray(key_bytes, dtype=np.uint8).tobytes()
    for i in range(len(part)):
        h.update(b"%d\t%d\t" % (int(part[i]), int(count[i])) + kb[int(key_off[i]):int(key_off[i + 1])] + b"\n")
    return h.hexdigest()


def assert_rows_equal(got, want, what=""):
    """got: _native.Rows (GPU); want: oracle.Rows.  Bit-exact: partitions, counts, offsets, key bytes."""
    assert got.n == want.n, f"{what}: {got.n} rows, oracle has {want.n}"
    if want.n == 0:
        return
    if not np.array_equal(got.key_off, want.key_off) or not np.array_equal(got.key_bytes, want.key_bytes):
        gk, wk = got.keys(), want.keys()
        for i, (a, b) in enumerate(zip(gk, wk)):
            assert a == b, f"{what}: row {i}: key {a!r} != oracle {b!r} (parts {got.part[i]} / {want.part[i]})"
    assert np.array_equal(got.part, want.part), f"{what}: partition ids differ"
    bad = np.nonzero(got.count != want.count)[0]
    assert bad.size == 0, f"{what}: {bad.size} counts differ, first at row {bad[0]}: " \
                          f"{got.count[bad[0]]} != {want.count[bad[0]]} key {want.key(int(bad[0]))!r}"


def sort_word(key):
    """(first 7 key bytes big-endian) << 8 | min(len, 8): the sort word of include/mrgpu.h
    (mr_shuffle_set_owners), restated for the tests."""
    w = 0
    for i in range(7):
        w = (w << 8) | (key[i] if i < len(key) else 0)
    return (w << 8) | min(len(key), 8)


def merge_iterator(runs):
    """utils.merge_iterator (utils.lua:206-271) restated: `runs` are iterables of (key, values) in
    ascending key order; a heap holds one line per run ordered by key (utils.lua:214), the smallest is
    popped and the value lists of all equal keys are concatenated (merge_min_keys, utils.lua:232-247).
    Yields (key, values) in ascending key order."""
    import heapq
    its = [iter(r) for r in runs]
    heap = []
    for which, it in enumerate(its):
        for key, values in it:
            heapq.heappush(heap, (key, which, list(values)))
            break
    while heap:
        key, which, values = heapq.heappop(heap)
        for k2, v2 in its[which]:
            heapq.heappush(heap, (k2, which, list(v2)))
            break
        while heap and heap[0][0] == key:
            _, w2, v2 = heapq.heappop(heap)
            values += v2
            for k3, v3 in its[w2]:
                heapq.heappush(heap, (k3, w2, list(v3)))
                break
        yield key, values
