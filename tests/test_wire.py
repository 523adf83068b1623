"""The reference's text wire format (utils.lua:100-120, job.lua:208-214,272-273) -- CPU only."""
from lua_mapreduce_1_b200 import corpus, wire


def test_escape_known_answers_from_the_reference():
    # utils.lua:345-348
    assert wire.escape(120) == b"120"
    assert wire.escape("30") == b'"30"'
    assert wire.escape("30\n") == b'"30\\n"'
    assert wire.serialize_table_ipairs([1, 2, 3, "hola"]) == b'{1,2,3,"hola"}'


def test_lua_q_rules():
    assert wire.lua_quote(b'a"b\\c') == b'"a\\"b\\\\c"'
    assert wire.lua_quote(b"a\x00b") == b'"a\\0b"' and wire.lua_quote(b"a\x001") == b'"a\\0001"'
    assert wire.lua_quote(b"\r\t") == b'"\\13\\9"' and wire.lua_quote(b"\x7f") == b'"\\127"'
    assert wire.lua_quote(b"caf\xc3\xa9") == b'"caf\xc3\xa9"'          # bytes >= 0x80 pass through
    assert wire.lua_number(5000) == "5000" and wire.lua_number(1e15) == "1e+15"   # %.14g, App. C-6
    assert wire.lua_number(2.5) == "2.5"


def test_line_roundtrip_on_torture_keys():
    keys = set(corpus.TORTURE.split()) | {b'q"uote', b"back\\slash", b"nl\nin", b"\x01\x02 3".replace(b" ", b"9")}
    for k in keys:
        line = wire.format_line(k, [7])
        assert b"\n" not in line[:-1]                                    # one chunk per line
        assert wire.parse_line(line) == (k, [7])
    assert wire.format_line(b"the", [85]) == b'return "the",{85}\n'
    assert wire.parse_line(b'return 3,{1,1}') == (3, [1, 1])             # numeric keys, utils.lua:360-362
    assert wire.parse_line(b'return "a",{1,"x",2.5}') == (b"a", [1, b"x", 2.5])


def test_file_names():
    assert wire.result_name("result", 3, 14) == "result.P03" and wire.result_name("result", 3, 9) == "result.P3"
    assert wire.map_result_name("map_results", 12, "4") == "map_results.P12.M4"
    assert sorted(wire.result_name("r", p, 14) for p in (10, 2, 1)) == ["r.P01", "r.P02", "r.P10"]


def test_merge_iterator_restatement_gives_the_reference_known_answer(tmp_path):
    # utils.lua:360-380: f1 = {1:{1,1}, 2:{1}, 3:{1}}, f2 = {1:{1,1,1,1}, 3:{1}, 4:{1}}
    import os
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from helpers import merge_iterator
    f1 = [(1, [1, 1]), (2, [1]), (3, [1])]
    f2 = [(1, [1, 1, 1, 1]), (3, [1]), (4, [1])]
    assert list(merge_iterator([f1, f2])) == [(1, [1] * 6), (2, [1]), (3, [1, 1]), (4, [1])]
    # and through files in the run format
    for name, run in (("map_results.P0.M1", f1), ("map_results.P0.M2", f2)):
        with open(tmp_path / name, "wb") as f:
            for k, v in run:
                f.write(wire.format_line(k, v))
    runs = [wire.read_run(str(tmp_path / n)) for n in wire.list_map_results(str(tmp_path), 0)]
    assert list(merge_iterator(runs)) == [(1, [1] * 6), (2, [1]), (3, [1, 1]), (4, [1])]
    assert wire.partitions_with_runs(str(tmp_path)) == [0]
