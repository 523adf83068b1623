"""BASELINE.json configs 3, 4 and 5 at the sizes it states (one GPU's share where the config names 8
GPUs), bit-exact against the oracle, inside the -m gpu suite.  Each test also appends its JSON line to
gpurun_out/configs_3_4_5.jsonl (when that directory exists) -- the evidence kept under profiles/.

  config 3  the 10x corpus: 1,970 splits, ~491 M tokens, word count (sharded over 2/4/8 ranks in
            bench.py's strong_config3 block; here all of it on one GPU)
  config 4  bigram count with ~5 x 10^7 distinct keys: hash-table overflow into the global table, key
            pool, heavy tie groups in the sort
  config 5  integer-key group-by-sum, Zipf(1.2) keys with four 6 % heavy hitters: one GPU's eighth of
            the 10^9 records (the 8-GPU run is bench-side: tests/mp_config5.py under torchrun)
"""
import json
import os

import pytest

import run_config

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _record(out):
    d = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(d):
        with open(os.path.join(d, "configs_3_4_5.jsonl"), "a") as f:
            f.write(json.dumps(out) + "\n")


@pytest.mark.timeout(900)
def test_config3_ten_times_corpus_full_size():
    out = run_config.run(3)
    _record(out)
    assert out["bit_exact"] and out["sum_ok"]
    assert out["splits"] == 1970 and out["tokens"] == out["emits"] > 490_000_000


@pytest.mark.timeout(1500)
def test_config4_bigrams_fifty_million_distinct_keys():
    out = run_config.run(4)
    _record(out)
    assert out["bit_exact"] and out["sum_ok"]
    assert out["distinct_keys"] > 40_000_000, out["distinct_keys"]


@pytest.mark.timeout(900)
def test_config5_one_gpus_share_of_a_billion_records():
    out = run_config.run(5)
    _record(out)
    assert out["bit_exact"] and out["sum_ok"]
    assert out["records"] == 125_000_000 == out["emits"]
