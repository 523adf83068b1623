"""A small run of every kernel family for compute-sanitizer (word count with tie groups and long keys,
bigrams, group-by, the no-allocation entry point, multi-round staging)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from lua_mapreduce_1_b200 import _native, corpus  # noqa: E402
from oracle import oracle as O  # noqa: E402


def same(got, want):
    return (got.n == want.n and np.array_equal(got.part, want.part) and np.array_equal(got.count, want.count)
            and np.array_equal(got.key_bytes, want.key_bytes))


splits = [corpus.generate_split(1, 300)[0].tobytes(), corpus.TORTURE,
          b" ".join(b"prefix_" + bytes([97 + i % 3, 97 + i % 5, 97 + i % 7]) * (1 + i % 4) for i in range(400))]
ctx = _native.Context(10, 1 << 20, max_keys=1 << 14, table_slots=1 << 15)
for i, s in enumerate(splits):
    ctx.map_split(i, s)
    if i == 0:
        ctx.map_finish()
assert same(ctx.rows(), O.run_naive(splits, 10))
ctx.job_begin(keep_input=True)
assert same(ctx.rows(), O.run_naive(splits, 10))
ctx.close()

ctx = _native.Context(10, 1 << 20, max_keys=1 << 15, table_slots=1 << 16, map_kind=_native.MR_MAP_BIGRAM_WS)
for i, s in enumerate(splits):
    ctx.map_split(i, s)
assert same(ctx.rows(), O.run_bigram(splits, 10))
ctx.close()

rng = np.random.default_rng(1)
k, v = rng.integers(0, 3000, 50000).astype(np.uint32), rng.integers(0, 9, 50000).astype(np.uint32)
ctx = _native.Context(7, 1 << 20, max_keys=1 << 13, table_slots=1 << 14, map_kind=_native.MR_MAP_RECORDS_U32,
                      partition_kind=_native.MR_PARTITION_MOD_U32)
ctx.map_records(0, k, v)
assert same(ctx.rows(), O.groupby_u32([(k, v)], 7))
ctx.close()

import torch  # noqa: E402
text = torch.from_numpy(np.frombuffer(splits[0], dtype=np.uint8).copy()).cuda()
out = torch.zeros(24 * 8192, dtype=torch.uint8, device="cuda")
n_out = torch.zeros(1, dtype=torch.int64, device="cuda")
_native.count_tokens_static(text.data_ptr(), text.numel(), out.data_ptr(), 8192, n_out.data_ptr())
print("sanitize_small ok", int(n_out.item()))
