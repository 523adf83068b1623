"""The C worker of examples/c_host (no Python, no torch on the product side) against the oracle:
result.P<p> files with "count\\tkey" lines, partition by partition, in the reference's key order.
One rank always; one rank per GPU (the library's own NCCL transport, rendezvous through a file)
when the box has several."""
import os
import subprocess

import numpy as np
import pytest

from lua_mapreduce_1_b200 import build as B, corpus
from oracle import oracle as O

pytestmark = pytest.mark.gpu
R = 10


def want_files(splits):
    rows = O.run_naive(splits, R)
    keys = rows.keys()
    out = {}
    for p, k, c in zip(rows.part, keys, rows.count):
        out.setdefault(int(p), []).append(b"%d\t%s\n" % (int(c), k))
    return {p: b"".join(v) for p, v in out.items()}, rows.tokens


def run_hosts(tmp_path, world, paths):
    exe = B.build_c_host()
    out = tmp_path / f"out{world}"
    out.mkdir()
    idfile = str(tmp_path / f"nccl_id_{world}")
    procs = [subprocess.Popen([exe, str(r), str(world), str(R), idfile, str(out)] + paths,
                              stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True) for r in range(world)]
    logs = []
    for p in procs:
        try:
            o, e = p.communicate(timeout=300)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        logs.append((p.returncode, o, e))
    assert all(rc == 0 for rc, _, _ in logs), logs
    return out, logs


@pytest.fixture
def split_files(tmp_path):
    splits, _ = corpus.generate(lines=[1200, 1, 0, 2500, 800, 1700, 950], threads=4)
    splits.append(np.frombuffer(corpus.TORTURE, dtype=np.uint8))
    paths = []
    for i, s in enumerate(splits):
        p = tmp_path / f"split{i:03d}.txt"
        p.write_bytes(bytes(s))
        paths.append(str(p))
    return splits, paths


def check(out, splits, logs):
    want, tokens = want_files(splits)
    got = {int(f.split(".P")[1]): open(os.path.join(out, f), "rb").read() for f in os.listdir(out)}
    assert sorted(got) == sorted(want)          # no file for an empty partition
    for p in want:
        assert got[p] == want[p], f"partition {p} differs"
    assert sum(int(o.split(":")[1].split()[0]) for _, o, _ in logs) == tokens


def test_c_host_single_rank(tmp_path, split_files):
    splits, paths = split_files
    out, logs = run_hosts(tmp_path, 1, paths)
    check(out, splits, logs)


def test_c_host_one_rank_per_gpu(tmp_path, split_files):
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    splits, paths = split_files
    out, logs = run_hosts(tmp_path, n, paths)
    check(out, splits, logs)
