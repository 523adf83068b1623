"""Bigram count (BASELINE config 4): the WordCount task module with a mapfn that emits neighbouring
word pairs of a line.  Everything but mapfn and the `native` map kind is WordCount's."""
from ..WordCount import (FILES, associative_reducer, combinerfn, commutative_reducer, finalfn,  # noqa: F401
                         idempotent_reducer, partitionfn, reducefn, taskfn)
from .. import WordCount as _wc

native = {"map": "bigram_ws", "reduce": "sum_u64", "partition": "fnv1_lua52_double",
          "num_reducers": _wc.NUM_REDUCERS}


def init(arg=None):
    _wc.init(arg)
    native["num_reducers"] = _wc.NUM_REDUCERS


def mapfn(key, value, emit):
    with open(value, "rb") as f:
        for line in f.read().split(b"\n"):     # io.lines cuts at "\n" only
            prev = None
            for w in line.split():
                if prev is not None:
                    emit(prev + b" " + w, 1)
                prev = w
