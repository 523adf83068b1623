"""World-size-2 (and 3) runs of the shuffle host logic on CPU with the gloo backend."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT


@pytest.mark.parametrize("world", [2, 3])
def test_shuffle_host_logic_gloo(world):
    env = dict(os.environ, OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29611 + world),
           os.path.join(ROOT, "tests", "mp_shuffle_worker.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=240, env=env, cwd=ROOT)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-4000:]
    assert p.stdout.count(": ok") == world


def test_shard_splits_round_robin():
    from lua_mapreduce_1_b200 import parallel
    assert parallel.shard_splits(7, 3, 0) == [0, 3, 6] and parallel.shard_splits(7, 3, 2) == [2, 5]
    assert parallel.shard_splits(2, 4, 3) == []
    assert [parallel.owner_of_partition(p, 8) for p in range(10)] == [0, 1, 2, 3, 4, 5, 6, 7, 0, 1]


def test_split_rows_cuts_the_output_order_into_equal_shares():
    # default ownership, 3 partitions on 2 ranks: rank 0 holds P0 and P2, rank 1 holds P1
    from lua_mapreduce_1_b200 import parallel
    counts = [[10, 0, 6], [0, 4, 0]]
    assert parallel.split_rows(counts, 2) == [(1, 0)]            # position 10 of 20 = first row of P1 on rank 1
    assert parallel.split_rows(counts, 4) == [(0, 5), (1, 0), (0, 11)]   # positions 5, 10, 15 (P2 starts at row 10 of rank 0)
    # range ownership already in place: slices of one partition follow each other in rank order
    counts = [[5, 0], [3, 2], [0, 6]]
    assert parallel.split_rows(counts, 2) == [(1, 3)]            # position 8 of 16: P0 has 8 rows, so rank 1's first P1 row (its row 3)
    assert parallel.split_rows(counts, 4) == [(0, 4), (1, 3), (2, 2)]
    assert parallel.split_rows([[0, 0], [0, 0]], 2) == []
