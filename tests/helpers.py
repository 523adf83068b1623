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
