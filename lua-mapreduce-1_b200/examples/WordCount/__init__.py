"""WordCount task module, all roles in one module (reference: examples/WordCount/init.lua).

The functions are the semantic definition of the job, exactly as in the reference; the
`native` descriptor declares that the GPU data plane may run it without interpreting them.
"""
import math

NUM_REDUCERS = 15            # init.lua:2
FNV_prime = 16777619.0       # init.lua:3
offset_basis = 2166136261.0  # init.lua:4
MAX = 2.0 ** 32              # init.lua:5

native = {"map": "tokenize_ws", "reduce": "sum_u64", "partition": "fnv1_lua52_double",
          "num_reducers": NUM_REDUCERS}

FILES = []   # set through init(args): the split files (init.lua:11-16 hard-codes four)


def init(arg=None):
    if arg and arg.get("files") is not None:
        FILES[:] = arg["files"]
    if arg and arg.get("num_reducers"):
        global NUM_REDUCERS
        NUM_REDUCERS = int(arg["num_reducers"])
        native["num_reducers"] = NUM_REDUCERS


def taskfn(emit):
    for i, f in enumerate(FILES, 1):
        emit(i, f)


def mapfn(key, value, emit):
    with open(value, "rb") as f:
        for line in f:
            for w in line.split():      # bytes.split(): runs of ASCII whitespace, like [^%s]+
                emit(w, 1)


def partitionfn(key):
    h = offset_basis                     # Lua numbers are doubles: keep the float arithmetic
    for c in key:
        h = (h * FNV_prime) % MAX
        h = float(int(h) ^ c)
    return int(h - math.floor(h / NUM_REDUCERS) * NUM_REDUCERS)


def reducefn(key, values, emit):
    count = 0
    for v in values:
        count += v
    emit(count)


combinerfn = reducefn


def finalfn(pairs_iterator):
    for key, values in pairs_iterator:
        print(f"{values[0]}\t{key.decode('latin-1')}")
    return True


associative_reducer = True
commutative_reducer = True
idempotent_reducer = True
