"""Writes tests/golden/lib_rs_scenarios.json.

The reference is a Rust crate that cannot be built or imported in this container (no cargo), so these
vectors are a HAND TRANSCRIPTION of the literal input tables and literal expected output batches of the
reference's own end-to-end tests in /root/reference/src/lib.rs (the same expectations are asserted there
for stock DataFusion HashJoinExec and for all ten ParallelHashJoin versions).  Each scenario cites the
lines it was transcribed from, and tests/test_golden_extraction.py re-derives every input table, every
expected batch and the SQL of every scenario mechanically from the reference source wherever that tree is
present and compares them with this transcription.  Run:  python tests/golden/make_lib_rs_golden.py
"""
import json
import os

L3 = ["left"] * 3
R3 = ["right"] * 3

scenarios = [
    {
        "name": "inner_join_with_nulls", "ref": "src/lib.rs:149-193", "join_type": "inner",
        "left": {"id": [1, 2, None], "value": L3},
        "right": {"id": [None, 2, 3], "value": R3},
        "on": [["id", "id"]], "filter": None,
        "expected": [[2], ["left"], [2], ["right"]],
    },
    {
        "name": "inner_join_without_matches", "ref": "src/lib.rs:210-246", "join_type": "inner",
        "left": {"id": [1, 2, None], "value": L3},
        "right": {"id": [None, 4, 5], "value": R3},
        "on": [["id", "id"]], "filter": None,
        "expected": [[], [], [], []],
    },
    {
        "name": "left_join", "ref": "src/lib.rs:263-307", "join_type": "left",
        "left": {"id": [1, 2, None], "value": L3},
        "right": {"id": [1, 1, None], "value": R3},
        "on": [["id", "id"]], "filter": None,
        "expected": [[None, 1, 1, 2], ["left", "left", "left", "left"],
                     [None, 1, 1, None], [None, "right", "right", None]],
    },
    {
        "name": "left_semi_join", "ref": "src/lib.rs:324-372", "join_type": "left_semi",
        "left": {"id": [1, 2, None], "value": L3},
        "right": {"id": [1, 1, None], "value": R3},
        "on": [["id", "id"]], "filter": None,
        "expected": [[1], ["left"]],
    },
    {
        "name": "left_anti_join", "ref": "src/lib.rs:389-437", "join_type": "left_anti",
        "left": {"id": [1, 2, None], "value": L3},
        "right": {"id": [1, 1, None], "value": R3},
        "on": [["id", "id"]], "filter": None,
        "expected": [[None, 2], ["left", "left"]],
    },
    {
        "name": "right_join", "ref": "src/lib.rs:454-498", "join_type": "right",
        "left": {"id": [1, 1, None], "value": L3},
        "right": {"id": [1, 2, None], "value": R3},
        "on": [["id", "id"]], "filter": None,
        # the SQL adds `WHERE right.id IS NOT NULL`, applied after the join (lib.rs:478)
        "post_filter": "r.id is not null",
        "expected": [[1, 1, None], ["left", "left", None], [1, 1, 2], ["right", "right", "right"]],
    },
    {
        "name": "right_anti_join", "ref": "src/lib.rs:515-574", "join_type": "right_anti",
        # `right` (4 rows) is the probe side of the RightAnti plan, `left` the build side (lib.rs:549-557)
        "left": {"id": [1, 2, None], "value": L3},
        "right": {"id": [1, 2, 3, None], "value": ["right"] * 4},
        "on": [["id", "id"]], "filter": None,
        "expected": [[3, None], ["right", "right"]],
    },
    {
        "name": "full_join", "ref": "src/lib.rs:591-634", "join_type": "full",
        "left": {"id": [1, 2, None], "value": L3},
        "right": {"id": [2, 3, None], "value": R3},
        "on": [["id", "id"]], "filter": None,
        "expected": [[1, 2, None, None, None], ["left", "left", None, None, "left"],
                     [None, 2, 3, None, None], [None, "right", "right", "right", None]],
    },
    {
        "name": "full_join_with_filter", "ref": "src/lib.rs:651-726", "join_type": "full",
        "left": {"id": [1, 2, 3, None], "value": ["left", "left", "same", "left"]},
        "right": {"id": [2, 3, 4, None], "value": ["right", "same", "right", "right"]},
        "on": [["id", "id"]], "filter": "l.value != r.value",
        "expected": [[1, 2, 3, None, None, None, None],
                     ["left", "left", "same", None, None, None, "left"],
                     [None, 2, None, 3, 4, None, None],
                     [None, "right", None, "same", "right", "right", None]],
    },
]

# inner_join_no_filter (src/lib.rs:67-132, tables :809-831): base_table = 16 batches x 64 rows with
# id1..id4 = 0..1023 and value "hello"; small_table_1..4 = id 0..1023, value "world"; four chained inner joins.
chain = {
    "name": "inner_join_no_filter", "ref": "src/lib.rs:67-132,809-831", "join_type": "inner",
    "batch_size": 64, "batches": 16,
    "base_table": {"id1": "range(0,1024)", "id2": "range(0,1024)", "id3": "range(0,1024)",
                   "id4": "range(0,1024)", "value": "const('hello',1024)"},
    "small_tables": 4,
    "small_table": {"id": "range(0,1024)", "value": "const('world',1024)"},
    "joins": [["id1", "id"], ["id2", "id"], ["id3", "id"], ["id4", "id"]],
    "expected": "13 columns: id1..id4 = 0..1023, 'hello', then 4 x (id = 0..1023, 'world'); 1024 rows",
}

# Known-answer vectors from the reference's unit tests (chain order, bitset, generator).
kats = {
    "chain_order": [
        {"ref": "src/utils/concurrent_queued_self_hash_join_map.rs:579-594",
         "inserts": [[1, 2, 2, 1, 1, 4], [4, 2, 1]],
         "get_all": {"1": [8, 4, 3, 0], "2": [7, 2, 1], "4": [6, 5]}},
        {"ref": "src/utils/concurrent_queued_self_hash_join_map.rs:603-616",
         "inserts": [[1, 1, 3, 1]], "get_all": {"1": [3, 1, 0]}},
        {"ref": "src/utils/concurrent_queued_self_hash_join_map.rs:621-644",
         "inserts": [[1, 2, 1]] * 4 + [[3, 2, 1]] * 4,
         "get_all": {"1": [23, 20, 17, 14, 11, 9, 8, 6, 5, 3, 2, 0]}},
        {"ref": "src/utils/partitioned_concurrent_self_hash_join_map.rs:395-413",
         "inserts": [[1, 2, 1, 2]], "get_all": {"1": [2, 0], "2": [3, 1]}},
        {"ref": "src/utils/partitioned_concurrent_self_hash_join_map.rs:415-434",
         "inserts": [[1, 2, 1, 2], [3, 2, 1, 3]], "get_all": {"1": [6, 2, 0], "2": [5, 3, 1]}},
        {"ref": "src/utils/concurrent_self_hash_join_map.rs:251-257", "inserts": [], "get_all": {"123": []}},
    ],
    "version10_build": {
        "ref": "src/operator/version10/build_implementation.rs:98-178",
        "batches": [[1, 2, 3], [2, 4, 5], [1, 6, 7]],
        "expected_reversed_lookup": {"1": [0, 6], "2": [1, 3], "3": [2], "4": [4], "5": [5], "6": [7], "7": [8]},
        "missing": 999,
    },
    "bitset": [
        {"ref": "src/utils/concurrent_bit_set.rs:136-145", "capacity": 64, "set": [[0, 1, 2, 5, 10, 31, 32, 63]]},
        {"ref": "src/utils/concurrent_bit_set.rs:147-156", "capacity": 256, "set": [[0, 1, 32, 64, 112, 196, 197, 255]]},
        {"ref": "src/utils/concurrent_bit_set.rs:158-168", "capacity": 83, "set": [[0, 1, 32, 64, 82]]},
        {"ref": "src/utils/concurrent_bit_set.rs:170-176", "capacity": 83,
         "set": [[0, 1, 32, 64, 82], [0, 10, 32, 74, 82]], "set_count": 7},
        {"ref": "src/utils/concurrent_bit_set.rs:178-189", "capacity": 83,
         "set_all_except": [0, 1, 4, 12, 54, 63, 64, 81, 82], "unset_indices": [0, 1, 4, 12, 54, 63, 64, 81, 82]},
    ],
    "exponential": {"ref": "src/api_utils.rs:51-71", "min": 0, "max": 10, "expected": [0, 0, 0, 0, 1, 2, 2, 3, 5, 7]},
}

out = {"scenarios": scenarios, "chain_scenario": chain, "kats": kats}
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib_rs_scenarios.json")
with open(path, "w") as f:
    json.dump(out, f, indent=1)
print("wrote", path)
