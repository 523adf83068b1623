"""Group-by-sum on number keys (BASELINE config 5), all roles in one task module.

The functions are the semantic definition (what the reference would run in Lua: a mapfn that
emits (key, value) numbers, `key % NUM_REDUCERS` as partitionfn, the WordCount sum as
reducer/combiner); the `native` descriptor lets the GPU data plane run it without interpreting them.
A map job's value is a pair of equal-length uint32 arrays (keys, values).
"""
NUM_REDUCERS = 10
BATCHES = []   # set through init(args): list of (keys, values) uint32 arrays

native = {"map": "records_u32", "reduce": "sum_u64", "partition": "mod_u32", "num_reducers": NUM_REDUCERS}


def init(arg=None):
    global NUM_REDUCERS
    if arg and arg.get("batches") is not None:
        BATCHES[:] = arg["batches"]
    if arg and arg.get("num_reducers"):
        NUM_REDUCERS = int(arg["num_reducers"])
        native["num_reducers"] = NUM_REDUCERS


def taskfn(emit):
    for i, b in enumerate(BATCHES, 1):
        emit(i, b)


def mapfn(key, value, emit):
    keys, values = value
    for k, v in zip(keys, values):
        emit(int(k), int(v))


def partitionfn(key):
    return key % NUM_REDUCERS


def reducefn(key, values, emit):
    emit(sum(values))


combinerfn = reducefn


def finalfn(pairs_iterator):
    for key, values in pairs_iterator:
        print(f"{values[0]}\t{key}")
    return True


associative_reducer = True
commutative_reducer = True
idempotent_reducer = True
