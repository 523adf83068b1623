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


@pytest.fixture(scope="session")
def reference_fixture():
    """The reference's in-tree WordCount input (taskfn.lua:8-11); only present in the build container."""
    paths = [os.path.join(REFERENCE, f) for f in FIXTURE_FILES]
    if not all(os.path.exists(p) for p in paths):
        pytest.skip("reference tree not present on this box")
    return [open(p, "rb").read() for p in paths]
