"""mr_count_tokens_static: the entry point for callers that cannot allocate device memory.

Some hosts have handed the whole GPU to their own memory pool, or sit in a pipeline where the
device-wide synchronisation of cudaMalloc/cudaFree is not acceptable; for them the library offers one
call that allocates nothing (the caller owns input and output, the hash table is a fixed array inside
the library).  This test drives it through the C ABI, twice in a row on different inputs -- the second
call finds the first call's table in the shared array and must clear it -- and checks every (token,
count) pair against the context-based entry points, exactly, and against a count made on the host
(bytes.split() on the six whitespace bytes of mapfn.lua:5, and the CPU oracle), so that a fault shared
by the two CUDA paths cannot hide.  Collected last on purpose.
"""
import collections

import numpy as np
import pytest

from lua_mapreduce_1_b200 import _native, corpus

pytestmark = pytest.mark.gpu


def reference_counts(text):
    """The existing entry points: mr_map_split -> mr_reduce_all -> mr_rows_fetch."""
    ctx = _native.Context(10, 1 << 20, max_keys=1 << 17)
    try:
        ctx.map_split(0, text)
        rows = ctx.rows()
        return sorted(zip(rows.keys(), (int(c) for c in rows.count))), int(ctx.job_sync().n_tokens)
    finally:
        ctx.close()


def static_counts(text, capacity=1 << 16):
    import torch
    t = torch.from_numpy(np.frombuffer(text, dtype=np.uint8).copy()).cuda()      # the caller's buffers
    out = torch.zeros(capacity * 24, dtype=torch.uint8, device="cuda")
    n_out = torch.full((1,), 12345, dtype=torch.int64, device="cuda")
    ntok = _native.count_tokens_static(t.data_ptr(), t.numel(), out.data_ptr(), capacity, n_out.data_ptr(),
                                       torch.cuda.current_stream().cuda_stream)
    n = int(n_out.item())
    rec = out.cpu().numpy()[: n * 24].view(_native.TOKEN_COUNT_DTYPE)
    got = []
    for r in rec:
        key = bytes(r["key"]).ljust(15, b"\0")[: int(r["len"])]   # numpy strips trailing NULs of S15
        got.append((key, int(r["count"])))
    return sorted(got), ntok


def test_static_entry_point_matches_context_api_twice_in_a_row():
    first = corpus.generate_split(7, 700)[0].tobytes()           # ~105 KB, tokens of 1..14 bytes
    second = (b"  two\tcalls share one table \n" * 50 + corpus.generate_split(8, 400)[0].tobytes()
              + b" caf\xc3\xa9 a\x00b fifteen-bytes-x x end-no-newline")
    assert len(first) <= _native.STATIC_MAX_BYTES and len(second) <= _native.STATIC_MAX_BYTES
    from oracle import oracle as O
    for text in (first, second):
        want, want_tokens = reference_counts(text)
        got, got_tokens = static_counts(text)
        assert got_tokens == want_tokens
        assert got == want
        host = collections.Counter(text.split())          # bytes.split(): runs of the same six whitespace bytes
        assert got == sorted(host.items()) and got_tokens == sum(host.values())
        naive = O.run_naive([text], 10)
        assert got == sorted(zip(naive.keys(), (int(c) for c in naive.count))) and got_tokens == naive.tokens


def test_static_entry_point_rejects_what_it_cannot_hold():
    import torch
    big = torch.full((_native.STATIC_MAX_BYTES + 1,), 0x61, dtype=torch.uint8, device="cuda")
    out = torch.zeros(24 * 16, dtype=torch.uint8, device="cuda")
    n_out = torch.zeros(1, dtype=torch.int64, device="cuda")
    with pytest.raises(_native.NativeError) as e:
        _native.count_tokens_static(big.data_ptr(), big.numel(), out.data_ptr(), 16, n_out.data_ptr())
    assert e.value.code == _native.MR_ERR_INVALID and "MR_STATIC_MAX_BYTES" in str(e.value)
    with pytest.raises(_native.NativeError) as e:                  # a 16-byte token does not fit the inline key
        static_counts(b"ok " + b"x" * 16 + b" ok")
    assert e.value.code == _native.MR_ERR_OVERFLOW
    with pytest.raises(_native.NativeError) as e:                  # three distinct tokens, room for two
        static_counts(b"a b c a", capacity=2)
    assert e.value.code == _native.MR_ERR_NOMEM
    t = torch.from_numpy(np.frombuffer(b"a b c a", dtype=np.uint8).copy()).cuda()
    out2 = torch.zeros(2 * 24, dtype=torch.uint8, device="cuda")
    n2 = torch.zeros(1, dtype=torch.int64, device="cuda")
    with pytest.raises(_native.NativeError):
        _native.count_tokens_static(t.data_ptr(), t.numel(), out2.data_ptr(), 2, n2.data_ptr())
    assert int(n2.item()) == 3                                     # the capacity a retry needs
    assert static_counts(b"a b c a", capacity=3)[0] == [(b"a", 2), (b"b", 1), (b"c", 1)]   # and it recovers
