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
