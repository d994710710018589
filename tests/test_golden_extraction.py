"""The golden scenarios (tests/golden/lib_rs_scenarios.json) are a hand transcription of the reference's end-to-end tests.
This test re-derives them MECHANICALLY from the reference source where it is present (this container: /root/reference; the GPU
box has no reference tree and skips): every array literal of every cited test function of src/lib.rs is parsed -- the two
input tables and the expected output batch -- and compared with the transcription, value for value."""
import os
import re

import pytest

from helpers import load_golden

LIB_RS = "/root/reference/src/lib.rs"
pytestmark = pytest.mark.skipif(not os.path.exists(LIB_RS), reason="reference source tree not present on this machine")

ARRAY = re.compile(
    r'make_int_array(?:_it_nullable)?\(vec!\[(?P<ints>[^\]]*)\](?:\.into_iter\(\))?\)'
    r'|make_string_array(?:_nullable)?\(vec!\[(?P<strs>[^\]]*)\]\)'
    r'|make_string_constant_array\("(?P<cs>[^"]*)"\.to_string\(\),\s*(?P<cn>\d+)\)'
    r'|make_int_array_from_range\((?P<r0>\d+),\s*(?P<r1>\d+)\)')


def parse_arrays(text):
    out = []
    for m in ARRAY.finditer(text):
        if m.group("ints") is not None:
            vals = []
            for tok in [t.strip() for t in m.group("ints").split(",") if t.strip()]:
                vals.append(None if tok == "None" else int(re.fullmatch(r"(?:Some\()?(-?\d+)\)?", tok).group(1)))
            out.append(vals)
        elif m.group("strs") is not None:
            vals = []
            for tok in [t.strip() for t in m.group("strs").split(",") if t.strip()]:
                mm = re.fullmatch(r'(?:Some\()?"([^"]*)"\.to_string\(\)\)?', tok)
                vals.append(None if tok == "None" else mm.group(1))
            out.append(vals)
        elif m.group("cs") is not None:
            out.append([m.group("cs")] * int(m.group("cn")))
        else:
            out.append(list(range(int(m.group("r0")), int(m.group("r1")))))
    return out


def function_text(lines, ref):
    a, b = (int(x) for x in ref.split(":")[1].split("-"))
    return "\n".join(lines[a - 1:b])


def test_every_scenario_matches_the_reference_source():
    with open(LIB_RS) as f:
        lines = f.read().split("\n")
    golden = load_golden()
    checked = 0
    for sc in golden["scenarios"]:
        text = function_text(lines, sc["ref"])
        arrays = parse_arrays(text)
        want = [sc["left"][k] for k in sc["left"]] + [sc["right"][k] for k in sc["right"]] + list(sc["expected"])
        if re.search(r"assert_eq!\(total_length,\s*0\)", text):          # the reference states "no rows at all" instead of a batch
            assert all(len(c) == 0 for c in sc["expected"]), sc["name"]
            want = want[:len(arrays)]
        assert arrays == want, (sc["name"], arrays, want)
        checked += 1
    assert checked == len(golden["scenarios"]) >= 9


# what the SQL text of every cited test must contain for the transcribed join type / filter to be the right reading of it
SQL_MARKS = {
    "inner_join_with_nulls": [r"\bJOIN my_catalog\.my_schema\.right\b", r"ON left\.id = right\.id"],
    "inner_join_without_matches": [r"\bJOIN my_catalog\.my_schema\.right\b", r"ON left\.id = right\.id"],
    "left_join": [r"LEFT JOIN my_catalog\.my_schema\.right", r"ON left\.id = right\.id"],
    "left_semi_join": [r"WHERE EXISTS \(", r"WHERE left\.id = right\.id"],
    "left_anti_join": [r"WHERE NOT EXISTS \(", r"WHERE left\.id = right\.id"],
    "right_join": [r"FULL OUTER JOIN my_catalog\.my_schema\.right", r"WHERE right\.id IS NOT NULL"],      # the optimiser makes it a right join
    "right_anti_join": [r"WHERE NOT EXISTS \(", r"WHERE left\.id = right\.id"],
    "full_join": [r"FULL JOIN my_catalog\.my_schema\.right", r"ON left\.id = right\.id\""],
    "full_join_with_filter": [r"FULL JOIN my_catalog\.my_schema\.right", r"ON left\.id = right\.id AND left\.value != right\.value"],
}


def test_sql_of_every_scenario_says_what_the_transcription_says():
    with open(LIB_RS) as f:
        lines = f.read().split("\n")
    for sc in load_golden()["scenarios"]:
        text = function_text(lines, sc["ref"])
        assert re.search(r"async fn " + sc["name"] + r"\(", text), sc["name"]
        for pat in SQL_MARKS[sc["name"]]:
            assert re.search(pat, text), (sc["name"], pat)
        assert (sc.get("filter") is not None) == ("!=" in "".join(SQL_MARKS[sc["name"]])), sc["name"]
        assert (sc.get("post_filter") is not None) == ("IS NOT NULL" in "".join(SQL_MARKS[sc["name"]])), sc["name"]


def test_chain_scenario_matches_the_reference_source():
    """inner_join_no_filter (src/lib.rs:67-132) and register_tables (:796-820): the 4-join chain's tables and its 13 expected columns."""
    with open(LIB_RS) as f:
        src = f.read()
    lines = src.split("\n")
    ch = load_golden()["chain_scenario"]
    m = re.search(r"const BATCH_SIZE: i32 = (\d+);", src)
    assert m and int(m.group(1)) == ch["batch_size"]
    reg = src[src.index("fn register_tables"):src.index("fn make_catalog_from_schema")]
    assert [int(x) for x in re.findall(r"make_table\(\s*(\d+),", reg)] == [ch["batches"], ch["batches"]]
    assert re.search(r"for i in 1\.\.5", reg) and ch["small_tables"] == 4
    assert re.findall(r'\("(\w+)", DataType::(\w+)', reg) == [("id1", "Int32"), ("id2", "Int32"), ("id3", "Int32"), ("id4", "Int32"), ("value", "Utf8"),
                                                            ("id", "Int32"), ("value", "Utf8")]
    assert re.findall(r'make_string_constant_array\("(\w+)"', reg) == ["hello", "world"]
    n = ch["batch_size"] * ch["batches"]
    expected = parse_arrays(function_text(lines, "src/lib.rs:67-132"))
    rng, hello, world = list(range(n)), ["hello"] * n, ["world"] * n
    assert expected == [rng, rng, rng, rng, hello] + [rng, world] * 4
    sql = function_text(lines, "src/lib.rs:67-132")
    assert [(a, b) for a, b in re.findall(r"ON base_table\.(\w+) = small_table_\d\.(\w+)", sql)] == [tuple(j) for j in ch["joins"]]
