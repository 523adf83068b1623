"""C-ABI surface, corpus generator and host mirror -- no GPU needed (no compute calls)."""
import ctypes
import os
import re
import threading

import numpy as np
import pytest

from conftest import ROOT, _has_gpu
from lua_mapreduce_1_b200 import _native, corpus, mapreduce


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "mrgpu.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(mr_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(_native.EXPORTS)
    lib = _native.load()
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.mr_abi_version() == 1
    assert ctypes.sizeof(_native.Config) == 64


def test_partition_of_matches_known_answers():
    assert _native.partition_of(b"the", 15) == 8 and _native.partition_of(b"the", 10) == 3
    assert _native.partition_of(b"function", 15) == 3 and _native.partition_of(b"--", 10) == 6


@pytest.mark.skipif(_has_gpu(), reason="checks the no-GPU failure mode")
def test_no_gpu_fails_loudly():
    with pytest.raises(_native.NativeError) as e:
        _native.Context(10, 1 << 20)
    assert "no CPU path" in str(e.value)


def test_c_worker_builds_against_the_header_alone(tmp_path):
    # examples/c_host: gcc + include/mrgpu.h + libmrgpu.so, nothing else; without a GPU the
    # worker must stop with the library's own error text, not fall back to anything
    import subprocess
    from lua_mapreduce_1_b200 import build as B
    exe = B.build_c_host()
    p = subprocess.run([exe], capture_output=True, text=True)
    assert p.returncode == 2 and "usage" in p.stderr
    if not _has_gpu():
        f = tmp_path / "a.txt"
        f.write_text("a b a")
        p = subprocess.run([exe, "0", "1", "10", str(tmp_path / "id"), str(tmp_path), str(f)],
                           capture_output=True, text=True)
        assert p.returncode == 1 and "mr_ctx_create" in p.stderr and "no CPU path" in p.stderr


def test_lua_c_api_binding_type_checks_against_the_header():
    # bindings/lua_mrgpu.c is the Lua 5.2 C-API glue a maintainer of the reference would add; no Lua
    # here, so it is compiled (not linked) against include/mrgpu.h and the declarations of the Lua
    # functions it uses (tests/lua_stub, signatures of the 5.2 manual) with warnings as errors
    import subprocess
    p = subprocess.run(["gcc", "-std=gnu11", "-Wall", "-Wextra", "-Werror", "-fsyntax-only",
                        "-I", os.path.join(ROOT, "tests", "lua_stub"), "-I", os.path.join(ROOT, "include"),
                        os.path.join(ROOT, "bindings", "lua_mrgpu.c")], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr


def test_bad_config_is_rejected():
    lib = _native.load()
    h = ctypes.c_void_p()
    cfg = _native.Config(12, 0, 10, 0, 0, 0, 1 << 20, 0, 0, None, 1, 0)   # wrong struct_size
    assert lib.mr_ctx_create(ctypes.byref(cfg), ctypes.byref(h)) == _native.MR_ERR_INVALID
    cfg = _native.Config(ctypes.sizeof(_native.Config), 0, 10, 7, 0, 0, 1 << 20, 0, 0, None, 1, 0)  # unknown map kind
    assert lib.mr_ctx_create(ctypes.byref(cfg), ctypes.byref(h)) == _native.MR_ERR_UNSUPPORTED
    assert b"unsupported" in lib.mr_last_error(None)
    cfg = _native.Config(ctypes.sizeof(_native.Config), 0, 0, 0, 0, 0, 1 << 20, 0, 0, None, 1, 0)
    assert lib.mr_ctx_create(ctypes.byref(cfg), ctypes.byref(h)) == _native.MR_ERR_INVALID


def test_corpus_is_deterministic_and_shaped():
    a, ta = corpus.generate_split(3, 500)
    b, tb = corpus.generate_split(3, 500)
    c, _ = corpus.generate_split(4, 500)
    assert ta == tb and np.array_equal(a, b) and not np.array_equal(a[:64], c[:64])
    text = a.tobytes()
    assert text.count(b"\n") == 500 and text.endswith(b"\n") and b"  " not in text
    assert len(text.split()) == ta
    lines = corpus.split_lines()
    assert len(lines) == 197 and sum(lines) == 1_965_734 and lines[-1] == 5_734
    assert len(corpus.split_lines(10)) == 1970
    words = {corpus.word(t) for t in list(range(2000)) + list(range(2**19 - 2000, 2**19))}
    assert len(words) == 4000 and all(w.isalpha() and w.islower() for w in words)


def test_corpus_zipf_head():
    sp, tok = corpus.generate(lines=[4000] * 4)
    from collections import Counter
    cnt = Counter(w for s in sp for w in s.tobytes().split())
    assert sum(cnt.values()) == tok
    top = cnt.most_common(2)
    assert top[0][0] == corpus.word(0) and top[1][0] == corpus.word(1)
    assert 1.6 < top[0][1] / top[1][1] < 2.4     # Zipf s=1: rank 1 about twice rank 2


class _NoNative:
    @staticmethod
    def init(arg=None): pass
    @staticmethod
    def taskfn(emit): emit(1, b"a b a")
    @staticmethod
    def mapfn(key, value, emit): pass
    @staticmethod
    def partitionfn(key): return 0
    @staticmethod
    def reducefn(key, values, emit): pass
    @staticmethod
    def finalfn(it): return True


def test_server_configure_validates_like_the_reference():
    s = mapreduce.server.new("localhost", "t1")
    with pytest.raises(AssertionError):
        s.configure({"taskfn": _NoNative})                       # mandatory fields (server.lua:425)
    s.configure({r: _NoNative for r in ("taskfn", "mapfn", "partitionfn", "reducefn", "finalfn")})
    assert s.configured
    w = mapreduce.worker.new("localhost", "t1")
    with pytest.raises(AssertionError):
        w.configure({"bogus": 1})                                 # worker.lua:142-148


def test_module_without_native_descriptor_is_refused():
    s = mapreduce.server.new("localhost", "t2")
    s.configure({r: _NoNative for r in ("taskfn", "mapfn", "partitionfn", "reducefn", "finalfn")})
    w = mapreduce.worker.new("localhost", "t2")
    err = []
    t = threading.Thread(target=lambda: err.append(pytest.raises(NotImplementedError, w.execute)))
    th = threading.Thread(target=lambda: None)
    t.start()
    # the server would wait for ever for a worker that refuses; only check the worker side
    b = mapreduce._board("localhost", "t2")
    with b.lock:
        b.task = {"status": mapreduce.TASK_MAP, "iteration": 1,
                  "params": {"mods": s.mods, "init_args": None, "result_ns": "result"}}
        b.lock.notify_all()
    t.join(20)
    assert not t.is_alive() and err and "no CPU fallback" in str(err[0].value)
    mapreduce.utest()
