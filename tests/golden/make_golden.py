"""Regenerates tests/golden/*.json.  Run in the build container (needs /root/reference):

    python tests/golden/make_golden.py

wordcount_fixture.json  known answers for the reference's own WordCount input (the four
                        Lua files of examples/WordCount/taskfn.lua:8-11), computed by a
                        pure-Python restatement that is independent of oracle/mr_oracle.c.
wordcount_fixture/      that input as DATA for the GPU box (which has no reference tree): the four
                        files' bytes back to back, zlib-compressed, plus their lengths and sha256
                        (input.bin.z, manifest.json).  It is the one input the reference's own
                        test.sh runs (test.sh:11-15); it is read as bytes by the tests, never as code.
torture_rows.json       full expected rows for the torture split of corpus.TORTURE.
"""
import collections
import hashlib
import json
import math
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

WS = b" \t\n\x0b\x0c\r"


def tokens(data):
    # mapfn.lua:4-5: maximal runs of non-%s bytes
    out, cur = [], bytearray()
    for c in data:
        if c in WS:
            if cur:
                out.append(bytes(cur)); cur = bytearray()
        else:
            cur.append(c)
    if cur:
        out.append(bytes(cur))
    return out


def fnv_lua52(key):
    # partitionfn.lua:10-15 with Python floats (IEEE binary64) and Lua 5.2's float modulo
    h = 2166136261.0
    for c in key:
        p = h * 16777619.0
        h = p - math.floor(p / 4294967296.0) * 4294967296.0
        h = float(int(h) ^ c)
    return int(h)


def rows(splits, R):
    cnt = collections.Counter()
    for s in splits:
        cnt.update(tokens(s))
    out = [(fnv_lua52(k) % R, k, v) for k, v in cnt.items()]
    out.sort(key=lambda r: (r[0], r[1]))      # bytes compare = unsigned lexicographic, shorter first
    return out


def digest(rws):
    h = hashlib.sha256()
    for p, k, v in rws:
        h.update(b"%d\t%d\t" % (p, v) + k + b"\n")
    return h.hexdigest()


def main():
    ref = "/root/reference/"
    files = ["mapreduce/server.lua", "mapreduce/worker.lua", "mapreduce/test.lua", "mapreduce/utils.lua"]
    splits = [open(ref + f, "rb").read() for f in files]
    g = {"files": files, "bytes": sum(map(len, splits))}
    toks = [t for s in splits for t in tokens(s)]
    cnt = collections.Counter(toks)
    g["tokens"], g["distinct"] = len(toks), len(cnt)
    g["top"] = [[k.decode(), v] for k, v in sorted(cnt.items(), key=lambda kv: (-kv[1], kv[0]))[:12]]
    g["hash"] = {k: fnv_lua52(k.encode()) for k in ["the", "local", "end", "function", "=", "--"]}
    for R in (15, 10):
        r = rows(splits, R)
        per = [0] * R
        for p, _, _ in r:
            per[p] += 1
        g[f"R{R}"] = {"keys_per_partition": per, "sha256": digest(r),
                      "first": [[p, k.decode("latin-1"), v] for p, k, v in r[:5]]}
    json.dump(g, open(os.path.join(HERE, "wordcount_fixture.json"), "w"), indent=1)
    fx = os.path.join(HERE, "wordcount_fixture")
    os.makedirs(fx, exist_ok=True)
    import zlib
    open(os.path.join(fx, "input.bin.z"), "wb").write(zlib.compress(b"".join(splits), 9))
    json.dump({"files": files, "lengths": [len(s) for s in splits],
               "sha256": [hashlib.sha256(s).hexdigest() for s in splits]},
              open(os.path.join(fx, "manifest.json"), "w"), indent=1)

    from lua_mapreduce_1_b200 import corpus
    t = {}
    for R in (15, 10, 1):
        r = rows([corpus.TORTURE], R)
        t[f"R{R}"] = {"rows": [[p, k.hex(), v] for p, k, v in r], "sha256": digest(r)}
    json.dump(t, open(os.path.join(HERE, "torture_rows.json"), "w"), indent=0)
    print(g["tokens"], g["distinct"], g["R15"]["keys_per_partition"])


if __name__ == "__main__":
    main()
