"""SASS opcode histogram of the built library (cuobjdump -sass), whole library and per kernel:
    python scripts/sass_histogram.py > profiles/r02/sass_opcode_histogram.txt"""
import collections, hashlib, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "lua-mapreduce-1_b200", "libmrgpu.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
whole, per, cur = collections.Counter(), {}, None
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1); per[cur] = collections.Counter(); continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
    if m and cur:
        whole[m.group(1)] += 1; per[cur][m.group(1)] += 1
md5 = hashlib.md5(open(lib, "rb").read()).hexdigest()
print(f"SASS opcode histogram of lua-mapreduce-1_b200/libmrgpu.so (sm_100a cubins; cuobjdump -sass), md5 {md5}")
tma = [k for k in whole if k.startswith(("UTMA", "UTCMMA", "UTCHMMA", "HMMA", "IMMA"))]
print("TMA / tensor-core opcodes: " + (", ".join(tma) if tma else "none (no TMA, no tensor cores: nothing here is a contraction); memory ops are LDG/STG/LDS/STS/ATOMS/ATOMG/RED/REDG."))
print("whole library: " + ", ".join(f"{k} {v}" for k, v in whole.most_common(40)))
print()
for fn in sorted(per, key=lambda f: -sum(per[f].values())):
    c = per[fn]
    print(f"{fn[:110]:110s} {sum(c.values()):6d} instr | " + ", ".join(f"{k} {v}" for k, v in c.most_common(12)))
