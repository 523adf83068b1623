import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device here")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


REFERENCE = "/root/reference"
FIXTURE_FILES = ("mapreduce/server.lua", "mapreduce/worker.lua", "mapreduce/test.lua", "mapreduce/utils.lua")
FIXTURE_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "wordcount_fixture")


def load_wordcount_fixture():
    """The reference's in-tree WordCount input (taskfn.lua:8-11, the input of test.sh:11-15) as four
    byte strings.  Read from the committed data blob (tests/golden/wordcount_fixture/, written by
    make_golden.py), so it is the same on the GPU box, where no reference tree exists; where the tree
    is mounted the blob is checked against it."""
    import hashlib
    import json
    import zlib
    man = json.load(open(os.path.join(FIXTURE_DIR, "manifest.json")))
    blob = zlib.decompress(open(os.path.join(FIXTURE_DIR, "input.bin.z"), "rb").read())
    assert list(man["files"]) == list(FIXTURE_FILES) and len(blob) == sum(man["lengths"])
    out, at = [], 0
    for n, sha in zip(man["lengths"], man["sha256"]):
        out.append(blob[at:at + n])
        at += n
        assert hashlib.sha256(out[-1]).hexdigest() == sha
    paths = [os.path.join(REFERENCE, f) for f in FIXTURE_FILES]
    if all(os.path.exists(p) for p in paths):
        assert out == [open(p, "rb").read() for p in paths], "fixture blob is stale: rerun tests/golden/make_golden.py"
    return out


@pytest.fixture(scope="session")
def reference_fixture():
    return load_wordcount_fixture()
