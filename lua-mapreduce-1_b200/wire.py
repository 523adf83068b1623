"""The reference's text wire format for intermediate runs and results.

Every run / result file is a sequence of Lua chunks, one per line:

    return <key>,{<v1>,<v2>,...}\\n           job.lua:208-214 (runs), job.lua:272-273 (results)

with `escape` (utils.lua:100-112): numbers through tostring (= "%.14g" in Lua 5.2), strings
through string.format("%q") with backslash-newline rewritten to \\n; values through
serialize_table_ipairs (utils.lua:114-120).  Files are named
`<path>/map_results.P<part>.M<mapper>` (job.lua:208-221) and `result.P<zero-padded part>`
(server.lua:313-321).  Writing rows in this format lets the unmodified server_final
(server.lua:346-385: sort file names, load() every line) and finalfn consume GPU results, and
parsing it lets GPU reducers consume runs written by reference mappers.
Host-side text formatting only; nothing here is on the device data path.
"""
import os
import re


def lua_number(v):
    """Lua 5.2 tostring(number): LUA_NUMBER_FMT = "%.14g" (integers print without a point)."""
    return "%.14g" % v


def lua_quote(s: bytes) -> bytes:
    """string.format("%q", s) of Lua 5.2 (lstrlib.c addquoted) followed by escape()'s gsub."""
    out = bytearray(b'"')
    n = len(s)
    for i, c in enumerate(s):
        if c in (0x22, 0x5C):                      # " and \ : backslash + the character
            out += b"\\" + bytes([c])
        elif c == 0x0A:                            # %q gives backslash-newline; escape() makes it \n
            out += b"\\n"
        elif c == 0 or c < 0x20 or c == 0x7F:      # '\0' or iscntrl in the C locale
            nxt_digit = i + 1 < n and 0x30 <= s[i + 1] <= 0x39
            out += (b"\\%03d" if nxt_digit else b"\\%d") % c
        else:
            out.append(c)
    out += b'"'
    return bytes(out)


def escape(v) -> bytes:
    """utils.lua:100-112."""
    if isinstance(v, (int, float)) and not isinstance(v, bool):
        return lua_number(v).encode()
    if isinstance(v, str):
        v = v.encode("latin-1")
    return lua_quote(bytes(v))


def serialize_table_ipairs(values) -> bytes:
    """utils.lua:114-120."""
    return b"{" + b",".join(escape(v) for v in values) + b"}"


def format_line(key, values) -> bytes:
    return b"return " + escape(key) + b"," + serialize_table_ipairs(values) + b"\n"


_ESC = {ord("a"): 7, ord("b"): 8, ord("f"): 12, ord("n"): 10, ord("r"): 13, ord("t"): 9, ord("v"): 11,
        ord("\\"): 0x5C, ord('"'): 0x22, ord("'"): 0x27, 10: 10}


def _parse_string(b, i):
    assert b[i] == 0x22
    out = bytearray()
    i += 1
    while b[i] != 0x22:
        c = b[i]
        if c != 0x5C:
            out.append(c)
            i += 1
            continue
        c = b[i + 1]
        if 0x30 <= c <= 0x39:                      # \ddd, up to three decimal digits
            j = i + 1
            while j < i + 4 and j < len(b) and 0x30 <= b[j] <= 0x39:
                j += 1
            out.append(int(b[i + 1:j]))
            i = j
        elif c == ord("x"):
            out.append(int(b[i + 2:i + 4], 16))
            i += 4
        else:
            out.append(_ESC[c])
            i += 2
    return bytes(out), i + 1


_NUM = re.compile(rb"-?(?:\d+\.?\d*(?:[eE][-+]?\d+)?|\.\d+)")


def _parse_value(b, i):
    if b[i] == 0x22:
        return _parse_string(b, i)
    m = _NUM.match(b, i)
    assert m, f"cannot parse value at {i}: {b[i:i + 20]!r}"
    t = m.group()
    v = float(t)
    return (int(v) if v.is_integer() and abs(v) < 2 ** 53 else v), m.end()


def parse_line(line: bytes):
    """Inverse of format_line = what load(line)() yields in utils.lua:222-224 / server.lua:381."""
    b = line.rstrip(b"\n")
    assert b.startswith(b"return "), b[:20]
    key, i = _parse_value(b, 7)
    assert b[i:i + 2] == b",{", b[i:i + 8]
    i += 2
    values = []
    while b[i] != 0x7D:
        v, i = _parse_value(b, i)
        values.append(v)
        if b[i] == 0x2C:
            i += 1
    return key, values


def count_digits(n):
    """server.lua count_digits: decimal digits of the largest partition id."""
    return max(1, len(str(int(n))))


def result_name(result_ns, part, max_part):
    """server.lua:313-321."""
    return "%s.P%0*d" % (result_ns, count_digits(max_part), part)


def map_result_name(map_results_ns, part, mapper_key):
    """job.lua:208-209."""
    return "%s.P%d.M%s" % (map_results_ns, part, mapper_key)


def write_results(rows, directory, result_ns="result"):
    """Writes one `result.P<part>` file per non-empty partition (server.lua:297-324: partitions that
    received no key do not exist), keys ascending -- the `shared` storage layout of fs.lua.
    rows: _native.Rows.  Returns the file names in server_final's order."""
    os.makedirs(directory, exist_ok=True)
    parts = [p for p in range(len(rows.part_begin) - 1) if rows.part_begin[p + 1] > rows.part_begin[p]]
    if not parts:
        return []
    kb = rows.key_bytes.tobytes()
    names = []
    for p in parts:
        name = result_name(result_ns, p, parts[-1])
        with open(os.path.join(directory, name), "wb") as f:
            for i in range(int(rows.part_begin[p]), int(rows.part_begin[p + 1])):
                f.write(format_line(kb[int(rows.key_off[i]):int(rows.key_off[i + 1])], [int(rows.count[i])]))
        names.append(name)
    return sorted(names)


def read_pairs(directory, names):
    """server_final's pair iterator (server.lua:358-383) over files in name order."""
    for name in sorted(names):
        with open(os.path.join(directory, name), "rb") as f:
            for line in f.read().split(b"\n"):
                if line:
                    yield parse_line(line)


def write_map_results(rows, directory, mapper_key, map_results_ns="map_results"):
    """What a map job leaves for the reducers (job.lua:196-221): per partition that received a key,
    one run `<ns>.P<part>.M<mapper_key>`, keys ascending, one line `return <%q key>,{<value>}` per
    key (the combiner has already folded a key's values into one, job.lua:198-202).
    rows: _native.Rows of a world-size-1 context that mapped this mapper's splits (its rows ARE the
    combined map output, sorted by partition and key).  Returns the file names written."""
    os.makedirs(directory, exist_ok=True)
    kb = rows.key_bytes.tobytes()
    names = []
    for p in range(len(rows.part_begin) - 1):
        lo, hi = int(rows.part_begin[p]), int(rows.part_begin[p + 1])
        if hi == lo:
            continue
        name = map_result_name(map_results_ns, p, mapper_key)
        with open(os.path.join(directory, name), "wb") as f:
            for i in range(lo, hi):
                f.write(format_line(kb[int(rows.key_off[i]):int(rows.key_off[i + 1])], [int(rows.count[i])]))
        names.append(name)
    return names


def list_map_results(directory, part, map_results_ns="map_results"):
    """The runs of one partition (job.lua:253-260: file names matching ^<path>/map_results.P<p>\\..*)."""
    prefix = "%s.P%d." % (map_results_ns, part)
    if not os.path.isdir(directory):
        return []
    return sorted(n for n in os.listdir(directory) if n.startswith(prefix))


def partitions_with_runs(directory, map_results_ns="map_results"):
    """server_prepare_reduce (server.lua:294-312): the partitions that have at least one run."""
    pat = re.compile(r"^%s\.P(\d+)\.M.*$" % re.escape(map_results_ns))
    if not os.path.isdir(directory):
        return []
    return sorted({int(m.group(1)) for m in map(pat.match, os.listdir(directory)) if m})


def read_run(path):
    """One run as (key, values) pairs in file order (utils.lua:218-230: load(line)() per line)."""
    with open(path, "rb") as f:
        for line in f.read().split(b"\n"):
            if line:
                yield parse_line(line)


def ingest_runs(ctx, directory, part, map_results_ns="map_results", fold=sum, batch=1 << 16):
    """Feeds every run of partition `part` to a context through mr_map_pairs: the reduce job's read
    side (job.lua:253-265).  A line's value list ({1,1,1} from a mapper whose combiner did not run,
    utils.lua:242-245) is folded with `fold` (sum for the WordCount reducer) before it goes to the
    device.  Returns (runs read, pairs ingested)."""
    names = list_map_results(directory, part, map_results_ns)
    keys, vals, total = [], [], 0
    for name in names:
        for key, values in read_run(os.path.join(directory, name)):
            if not isinstance(key, bytes):
                raise TypeError("string keys only (number keys go through mr_map_records)")
            keys.append(key)
            vals.append(int(fold(values)))
            if len(keys) >= batch:
                ctx.map_pairs(total, keys, vals)
                total += len(keys)
                keys, vals = [], []
    if keys:
        ctx.map_pairs(total, keys, vals)
        total += len(keys)
    return len(names), total
