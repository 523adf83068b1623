"""Synthetic Europarl-shaped corpus.  Thin wrapper over csrc/corpus.c.

Shape of the reference's benchmark corpus (reference README.md:43-45): 1,965,734
lines in 197 splits of at most 10,000 lines.  `scale` repeats that shape.
"""
import ctypes as C
import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np

from . import build as _build

SEED_CONFIG2 = 0x5EED0001
EUROPARL_LINES = 1_965_734
LINES_PER_SPLIT = 10_000

_lib = None


def _load():
    global _lib
    if _lib is None:
        path = _build.build_corpus()
        L = C.CDLL(path)
        L.mrc_init.restype = C.c_int
        L.mrc_max_bytes.restype = C.c_uint64
        L.mrc_max_bytes.argtypes = [C.c_uint32]
        L.mrc_generate_split.restype = C.c_uint64
        L.mrc_generate_split.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_void_p,
                                         C.c_uint64, C.POINTER(C.c_uint64)]
        L.mrc_word.restype = C.c_uint32
        L.mrc_word.argtypes = [C.c_uint32, C.c_char_p]
        if L.mrc_init() != 0:
            raise MemoryError("corpus vocabulary allocation failed")
        _lib = L
    return _lib


def split_lines(scale=1):
    """Line count of every split: scale x (196 x 10,000 + 1 x 5,734)."""
    full, rest = divmod(EUROPARL_LINES, LINES_PER_SPLIT)
    one = [LINES_PER_SPLIT] * full + ([rest] if rest else [])
    return one * scale


def generate_split(split_id, lines, seed=SEED_CONFIG2):
    L = _load()
    cap = int(L.mrc_max_bytes(lines))
    buf = np.empty(cap, dtype=np.uint8)
    tokens = C.c_uint64(0)
    n = L.mrc_generate_split(seed, split_id, lines, buf.ctypes.data, cap, C.byref(tokens))
    if n == 0 and lines:
        raise RuntimeError("corpus buffer overflow")
    return buf[:n].copy(), int(tokens.value)


def generate(scale=1, seed=SEED_CONFIG2, lines=None, first_split=0, threads=None):
    """Returns (list of uint8 arrays, total tokens).  Split i always has the same bytes
    whatever subset is generated, so ranks can generate only their own shard."""
    lines = split_lines(scale) if lines is None else list(lines)
    ids = range(first_split, first_split + len(lines))
    threads = threads or min(32, os.cpu_count() or 1)
    with ThreadPoolExecutor(threads) as ex:
        out = list(ex.map(lambda a: generate_split(a[0], a[1], seed), zip(ids, lines)))
    return [o[0] for o in out], sum(o[1] for o in out)


def word(type_id):
    buf = C.create_string_buffer(16)
    n = _load().mrc_word(type_id, buf)
    return buf.raw[:n]


TORTURE = (
    b"  lead\ttabs\x0bvt\x0cff\rcr\r\nline two  double  spaces\n\n\n"
    b"quote\"in back\\slash caf\xc3\xa9 na\xc3\xafve \xa0nbsp\xa0 \xff\xfe\x80 "
    + b"L" * 300 + b" " + b"L" * 299 + b"M " + b"abcdefg abcdefgh abcdefghi abcdefg abcdefgh "
    + b"0123456789abcde 0123456789abcdef 0123456789abcdefg "
    + b"x" * 15 + b" " + b"x" * 16 + b" " + b"x" * 17 + b" " + b"x" * 31 + b" " + b"x" * 33
    + b"\nnul\x00in\x00side a\x00 a a\x00\x00 tail-no-newline"
)
