"""Parity of the CUDA data plane with the CPU oracle, through the C ABI.  Bit-exact:
same key set, same integer counts, same partition ids, same order (partition ascending,
key ascending in unsigned-byte order) -- SURVEY.md App. A."""
import random

import numpy as np
import pytest

from helpers import assert_rows_equal, golden, rows_digest
from lua_mapreduce_1_b200 import _native, corpus
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def run_gpu(splits, R, arena=None, max_keys=1 << 16, **kw):
    total = sum(len(s) for s in splits)
    ctx = _native.Context(R, arena or (total + 64 * len(splits) + 4096), max_keys=max_keys, **kw)
    try:
        for i, s in enumerate(splits):
            ctx.map_split(i, s)
        ctx.map_finish()
        ctx.reduce_all()
        info = ctx.job_sync()
        rows = ctx.rows()
        return rows, info, ctx.stats()
    finally:
        ctx.close()


def check(splits, R, **kw):
    want = O.run_naive(splits, R)
    got, info, st = run_gpu(splits, R, **kw)
    assert info.n_tokens == want.tokens
    assert info.input_bytes == sum(len(s) for s in splits)
    assert_rows_equal(got, want, f"R={R}")
    assert info.n_keys == want.n and info.key_bytes == len(want.key_bytes)
    pb = np.searchsorted(want.part, np.arange(R + 1), side="left")
    assert np.array_equal(got.part_begin.astype(np.int64), pb)
    return got, info


@pytest.mark.parametrize("text", [b"", b"   \n\t\r ", b"a", b"a ", b" a", b"ab cd ab", b"x" * 7, b"x" * 8,
                                  b"x" * 9, b"\n".join([b"w"] * 100), b"a\x00b a\x00b a"])
def test_tiny_inputs(text):
    check([text], 15)


def test_empty_job_and_empty_splits():
    check([], 10)
    check([b"", b"", b"one", b""], 10)


def test_torture_split_matches_oracle_and_golden():
    g = golden("torture_rows.json")
    for R in (15, 10, 1):
        got, _ = check([corpus.TORTURE], R)
        assert rows_digest(got.part, got.count, got.key_off, got.key_bytes) == g[f"R{R}"]["sha256"]


def test_tokens_never_span_splits():
    # "ab" ends split 0 without whitespace, "cd" starts split 1: must stay two tokens
    got, _ = check([b"x ab", b"cd y", b"ab"], 10)
    assert b"abcd" not in got.keys()


def test_alignment_and_segment_boundaries():
    # tokens straddling 16-byte, 32-byte and 1024-byte (segment) boundaries, every phase
    rnd = random.Random(11)
    for shift in range(0, 40):
        parts = [b" " * shift]
        while sum(map(len, parts)) < 5000:
            n = rnd.choice([1, 2, 3, 6, 7, 8, 9, 15, 16, 17, 31, 33, 70])
            parts.append(bytes(rnd.choice(b"abcxyz") for _ in range(n)))
            parts.append(rnd.choice([b" ", b"\n", b"  ", b"\t"]))
        check([b"".join(parts)], 10)


def test_long_token_spanning_many_segments():
    big = b"Q" * 5000
    check([b"a " + big + b" b " + big + b" " + big[:-1] + b" z"], 10)


def test_reference_shaped_fixture(reference_fixture):
    # config 1: the reference's own input (test.sh:11-15), from the committed data blob
    g = golden("wordcount_fixture.json")
    for R in (15, 10):
        got, info = check(reference_fixture, R)
        assert info.n_tokens == 4977 and info.n_keys == 1575
        assert rows_digest(got.part, got.count, got.key_off, got.key_bytes) == g[f"R{R}"]["sha256"]


def test_source_code_like_text():
    # stand-in for config 1 on boxes without the reference tree: punctuation-heavy, mixed lengths
    rnd = random.Random(5)
    vocab = [b"local", b"function", b"end", b"=", b"==", b"--", b"{", b"}", b"return", b"self.cnn:connect()",
             b"string.format(\"%s.P%d.M%s\",", b"if", b"then", b"utils.MAX_MAP_RESULT", b"job_mark_as_written(self,",
             b"--------------------------------------------------------------------------"]
    vocab += [bytes(rnd.choice(b"abcdefghijklmnopqrstuvwxyz_.():") for _ in range(rnd.randrange(1, 30)))
              for _ in range(400)]
    splits = []
    for _ in range(4):
        lines = []
        for _ in range(400):
            lines.append(b"  " * rnd.randrange(4) + b" ".join(rnd.choice(vocab) for _ in range(rnd.randrange(0, 9))))
        splits.append(b"\n".join(lines))
    for R in (15, 10):
        check(splits, R)


def test_many_keys_sharing_a_7_byte_prefix_small_path():
    rnd = random.Random(2)
    keys = {b"prefix_" + bytes(rnd.choice(b"abc") for _ in range(rnd.randrange(1, 12))) for _ in range(600)}
    keys |= {b"prefix_", b"prefix", b"prefix_a" + b"\x00", b"prefixes"}
    text = b" ".join(rnd.choice(sorted(keys)) for _ in range(4000))
    got, info = check([text], 10)
    assert info.tie_rounds == 1


def test_many_keys_sharing_a_7_byte_prefix_general_path():
    rnd = random.Random(4)
    keys = sorted({b"commonprefix/" + bytes(rnd.choice(b"abcdefgh") for _ in range(rnd.randrange(1, 20)))
                   for _ in range(6000)})
    text = b"\n".join(rnd.choice(keys) for _ in range(30000))
    got, info = check([text], 10, max_keys=1 << 15)
    assert info.tie_rounds == 2


def test_one_hundred_thousand_keys_in_large_tie_groups():
    # more than 64K keys in groups that agree in their first 7 bytes: the order comes from radix
    # rounds over the later bytes (run_lsd_ties), inline keys and keys longer than 15 bytes alike
    rnd = random.Random(9)
    keys = [b"shared/%06d" % i for i in range(60000)]
    keys += [b"sharedpath/%05d/%s" % (i, bytes(rnd.choice(b"xyz") for _ in range(rnd.randrange(1, 12)))) for i in range(50000)]
    keys += [b"other", b"shared/", b"shared"]
    rnd.shuffle(keys)
    text = b" ".join(keys) + b" " + b" ".join(keys[:5000])
    got, info = check([text], 10, max_keys=1 << 18)
    assert info.tie_rounds == 2 and info.n_keys == len(set(keys))


def test_synthetic_corpus_small():
    splits, tok = corpus.generate(lines=[3000, 3000, 1234, 1], first_split=190)
    got, info = check(splits, 10, max_keys=1 << 18)
    assert info.n_tokens == tok


def test_incremental_map_finish_and_keep_input():
    splits, tok = corpus.generate(lines=[500] * 6)
    want = O.run_staged(splits, 10, 2)
    ctx = _native.Context(10, 1 << 22, max_keys=1 << 17)
    try:
        for i, s in enumerate(splits):
            ctx.map_split(i, s)
            if i % 2 == 1:
                ctx.map_finish()          # three partial map rounds
        ctx.reduce_all()
        assert_rows_equal(ctx.rows(), want, "incremental")
        for _ in range(2):                # "loop" iteration over the resident splits
            ctx.job_begin(keep_input=True)
            ctx.map_finish()
            ctx.reduce_all()
            info = ctx.job_sync()
            assert info.n_tokens == tok
            assert_rows_equal(ctx.rows(), want, "keep_input")
        ctx.job_begin()                   # fresh job on the same context, smaller input
        ctx.map_split(0, b"fresh start fresh")
        assert_rows_equal(ctx.rows(), O.run_naive([b"fresh start fresh"], 10), "fresh")
    finally:
        ctx.close()


def test_more_splits_than_the_gap_list_holds():
    # the pinned list of inter-split gaps is a ring of 32768 entries that the host only waits on
    # when it wraps: 70,000 tiny splits wrap it twice inside one job, a second job wraps it again
    words = [b"w%d" % (i % 257) for i in range(70000)]
    splits = [w + b" tail" if i % 3 else w for i, w in enumerate(words)]
    want = O.run_naive(splits, 10)
    ctx = _native.Context(10, 1 << 23, max_keys=1 << 16)
    try:
        for rep in range(2):
            ctx.job_begin()
            for i, s in enumerate(splits):
                ctx.map_split(i, s)
            ctx.reduce_all()
            assert_rows_equal(ctx.rows(), want, f"job {rep}")
    finally:
        ctx.close()


def test_partition_cursor_matches_rows():
    splits, _ = corpus.generate(lines=[800, 800])
    ctx = _native.Context(15, 1 << 21, max_keys=1 << 17)
    try:
        for i, s in enumerate(splits):
            ctx.map_split(i, s)
        rows = ctx.rows()
        keys = rows.keys()
        seen = 0
        for p in range(15):
            lo, hi = int(rows.part_begin[p]), int(rows.part_begin[p + 1])
            cur = list(ctx.partition_cursor(p))
            assert [k for k, _ in cur] == keys[lo:hi]
            assert [c for _, c in cur] == [int(c) for c in rows.count[lo:hi]]
            seen += len(cur)
        assert seen == rows.n
        with pytest.raises(_native.NativeError):
            list(ctx.partition_cursor(15))
    finally:
        ctx.close()


def test_capacity_errors_are_reported():
    splits, _ = corpus.generate(lines=[2000])
    with pytest.raises(_native.NativeError) as e:
        run_gpu(splits, 10, arena=1 << 12)
    assert e.value.code == _native.MR_ERR_NOMEM
    with pytest.raises(_native.NativeError) as e:
        run_gpu(splits, 10, max_keys=1 << 10, table_slots=1 << 11)
    assert e.value.code in (_native.MR_ERR_NOMEM, _native.MR_ERR_OVERFLOW)


def test_splits_named_by_path(tmp_path):
    # mr_map_split_path: files of very different sizes through the one page-locked buffer, an empty
    # file, a missing file
    splits, _ = corpus.generate(lines=[3000, 5, 0, 1200, 40000, 7])
    splits.append(np.frombuffer(corpus.TORTURE, dtype=np.uint8))
    paths = []
    for i, s in enumerate(splits):
        p = tmp_path / f"s{i}.txt"
        p.write_bytes(bytes(s))
        paths.append(str(p))
    want = O.run_naive(splits, 10)
    ctx = _native.Context(10, 1 << 23, max_keys=1 << 18)
    try:
        for rep in range(2):
            ctx.job_begin()
            for i, p in enumerate(paths):
                ctx.map_split_path(i, p)
            ctx.reduce_all()
            assert_rows_equal(ctx.rows(), want, f"by path, job {rep}")
        with pytest.raises(_native.NativeError) as e:
            ctx.map_split_path(99, str(tmp_path / "missing.txt"))
        assert e.value.code == _native.MR_ERR_INVALID and "cannot open" in str(e.value)
    finally:
        ctx.close()


def test_device_memory_input():
    import torch
    splits, _ = corpus.generate(lines=[1000])
    t = torch.from_numpy(splits[0]).cuda()
    ctx = _native.Context(10, 1 << 21, max_keys=1 << 17, stream=torch.cuda.current_stream().cuda_stream)
    try:
        ctx.map_split_ptr(0, t.data_ptr(), t.numel(), device=True)
        assert_rows_equal(ctx.rows(), O.run_naive(splits, 10), "device input")
    finally:
        ctx.close()


def test_config2_full_size_exact():
    # BASELINE.json configs[1]: 197 splits, ~49 M tokens, R = 10 -- full comparison with the oracle
    splits, tok = corpus.generate()
    want = O.run_staged(splits, 10, O.max_threads())
    total = sum(len(s) for s in splits)
    ctx = _native.Context(10, total + 64 * len(splits) + 4096, max_keys=1 << 20, table_slots=1 << 21)
    try:
        for i, s in enumerate(splits):
            ctx.map_split(i, s)
        ctx.map_finish()
        ctx.reduce_all()
        info = ctx.job_sync()
        got = ctx.rows()
        assert info.n_tokens == tok == want.tokens
        assert_rows_equal(got, want, "config 2")
        # size-independent properties
        assert int(got.count.sum()) == tok
        assert np.all(np.diff(got.part.astype(np.int64)) >= 0)
    finally:
        ctx.close()


def test_native_transport_single_rank():
    # mr_nccl_init / mr_shuffle_nccl with a world of one: the slot goes to this rank itself
    splits, tok = corpus.generate(lines=[700, 1, 0, 1300])
    splits.append(np.frombuffer(corpus.TORTURE, dtype=np.uint8))
    want = O.run_naive(splits, 10)
    ctx = _native.Context(10, 1 << 22, max_keys=1 << 17)
    try:
        with pytest.raises(_native.NativeError):
            ctx.shuffle_nccl()                       # no communicator yet
        ctx.nccl_init(_native.nccl_unique_id())
        for i, s in enumerate(splits):
            ctx.map_split(i, s)
        ctx.map_finish()
        slot = ctx.shuffle_nccl()
        ctx.reduce_all()
        assert_rows_equal(ctx.rows(), want, "native transport")
        ctx.job_begin(keep_input=True)
        ctx.map_finish()
        assert ctx.shuffle_nccl(slot) == slot        # no host synchronisation this time
        ctx.reduce_all()
        assert_rows_equal(ctx.rows(), want, "native transport, fixed slot")
    finally:
        ctx.close()


def test_a_rank_of_many_must_shuffle_before_it_reduces():
    ctx = _native.Context(10, 1 << 20, max_keys=1 << 12, world_size=2, rank=0)
    try:
        ctx.map_split(0, b"a b a")
        ctx.map_finish()
        with pytest.raises(_native.NativeError) as e:
            ctx.reduce_all()
        assert e.value.code == _native.MR_ERR_INVALID and "shuffle" in str(e.value)
    finally:
        ctx.close()


def test_p2p_shuffle_single_rank():
    # the peer-to-peer shuffle with a world of one: the window is this context's own memory
    splits, tok = corpus.generate(lines=[900, 1, 0, 1100])
    splits.append(np.frombuffer(corpus.TORTURE, dtype=np.uint8))
    want = O.run_naive(splits, 10)
    ctx = _native.Context(10, 1 << 22, max_keys=1 << 17)
    try:
        for i, s in enumerate(splits):
            ctx.map_split(i, s)
        ctx.map_finish()
        with pytest.raises(_native.NativeError):
            ctx.shuffle_p2p()                        # no window yet
        ctx.p2p_attach([ctx.p2p_setup(4 << 20)])
        for it in range(3):
            if it:
                ctx.job_begin(keep_input=True)
                ctx.map_finish()
            ctx.shuffle_p2p()
            ctx.reduce_all()
            assert_rows_equal(ctx.rows(), want, f"p2p job {it}")
        list(ctx.partition_cursor(3))                # cursors read the rows, not the window
        ctx.p2p_attach([ctx.p2p_setup(64)])          # too small: reported, not written past the slot
        ctx.job_begin(keep_input=True)
        ctx.map_finish()
        ctx.shuffle_p2p()
        ctx.reduce_all()
        with pytest.raises(_native.NativeError) as e:
            ctx.job_sync()
        assert e.value.code == _native.MR_ERR_OVERFLOW
    finally:
        ctx.close()


def test_p2p_shuffle_two_contexts_in_one_process():
    # one process driving two GPUs: the windows are reached through peer access, not IPC.  (Not with
    # both ranks on one GPU: the owner's kernel waits on the GPU for its peer and can starve it there.)
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs at least 2 GPUs")
    two = True
    splits, tok = corpus.generate(lines=[800, 1200, 1, 0, 950, 700, 1300])
    splits.append(np.frombuffer(corpus.TORTURE, dtype=np.uint8))
    want = O.run_naive(splits, 10)
    ctxs = [_native.Context(10, 1 << 22, max_keys=1 << 17, device=r if two else 0, world_size=2, rank=r) for r in range(2)]
    try:
        blobs = [c.p2p_setup(4 << 20) for c in ctxs]
        for c in ctxs:
            c.p2p_attach(blobs)
        for it in range(3):
            for r, c in enumerate(ctxs):
                if it == 0:
                    for i in range(r, len(splits), 2):
                        c.map_split(i, splits[i])
                else:
                    c.job_begin(keep_input=True)
                c.map_finish()
            if not two:
                torch.cuda.synchronize()             # ranks sharing one GPU: maps done before anybody waits in a kernel
            for c in ctxs:
                c.shuffle_p2p()                      # asynchronous: neither call waits for the other rank
            for c in ctxs:
                c.reduce_all()
            keys = want.keys()
            for r, c in enumerate(ctxs):
                got = c.rows()
                owned = (want.part % 2) == r
                assert got.keys() == [k for k, o in zip(keys, owned) if o], (it, r)
                assert np.array_equal(got.count, want.count[owned]) and np.array_equal(got.part, want.part[owned])
    finally:
        for c in ctxs:
            c.close()


def test_multi_gpu_shuffle_parity():
    # all visible GPUs (>= 2), one rank each, nccl all-to-all; skipped on a single-GPU box
    import os
    import subprocess
    import sys
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", "29655", os.path.join(root, "tests", "mp_gpu_parity.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=root)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-4000:]
    assert p.stdout.count("multi-GPU parity ok") == n


def test_multi_gpu_fuzz():
    # randomised cases through every multi-GPU path; skipped on a single-GPU box
    import os
    import subprocess
    import sys
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={min(n, 4)}",
           "--master-addr", "127.0.0.1", "--master-port", "29657", os.path.join(root, "tests", "mp_gpu_fuzz.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=500, cwd=root)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-4000:]
    assert p.stdout.count("multi-GPU fuzz ok") == min(n, 4)
