"""C5 as a repeatable test: the join trees of TPC-H Q3 / Q5 / Q9 on TPC-H-shaped data (the reference's
tpc harness and dbgen data are unavailable on this box: no cargo, no docker, no network -- DESIGN.md) through the operator
(build_push / build_finish / probe_batch / gather), every join re-run with pyarrow Acero and compared (row count + key
equality on every output row); plus a Q3-shaped join carrying Utf8 payload columns compared as full rows."""
import json
import os
import subprocess
import sys

import numpy as np
import pyarrow as pa
import pytest

from helpers import rows_multiset

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("query", ["q3", "q5", "q9"])
def test_tpch_join_trees_verified_against_acero(query):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "tpch_shape.py"), "--sf", "0.05", "--queries", query, "--verify",
                          "--chunk", "100000"], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [json.loads(l) for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1 and lines[0]["query"] == query and lines[0]["verified_vs_acero"] is True
    assert lines[0]["result_rows"] > 0
    n_joins = {"q3": 2, "q5": 4, "q9": 5}[query]
    assert len(lines[0]["joins"]) == n_joins
    if query != "q3":
        assert any(j["keys"] == 2 for j in lines[0]["joins"])          # the composite-key join of Q5 / Q9


def test_q3_shaped_join_with_utf8_payload_rows():
    """customer[BUILDING] x orders with c_name / o_comment Utf8 payload columns: output ROWS against Acero."""
    import datafusion_parallelism_b200 as d
    rng = np.random.default_rng(3)
    n_cust, n_ord = 7_500, 75_000
    cust = pa.table({"c_custkey": pa.array(np.arange(1, n_cust + 1, dtype=np.int64)),
                     "c_name": pa.array([f"Customer#{i:09d}" for i in range(1, n_cust + 1)]),
                     "c_mktsegment": pa.array(rng.integers(0, 5, n_cust).astype(np.int32))})
    ck = rng.integers(1, n_cust + 1, n_ord).astype(np.int64)
    ck = np.where(ck % 3 == 0, np.maximum(ck - 1, 1), ck)                  # a third of the customers have no orders
    orders = pa.table({"o_custkey": pa.array(ck), "o_orderkey": pa.array(np.arange(n_ord, dtype=np.int64) * 4 + 1),
                       "o_comment": pa.array([None if i % 17 == 0 else "c" * int(i % 23) + str(i) for i in range(n_ord)])})
    build = cust.filter(pa.compute.equal(cust.column("c_mktsegment"), 0))
    with d.HashJoin("inner", key_widths=(8,), build_payload_types=(d.T_UTF8, 4), probe_payload_types=(8, d.T_UTF8)) as j:
        j.build_push(0, [build.column("c_custkey")], [build.column("c_name"), build.column("c_mktsegment")])
        j.build_finish(0)
        pieces = []
        for lo in range(0, n_ord, 8192):                                    # DataFusion-sized batches
            pb = orders.slice(lo, 8192).combine_chunks()
            res = j.probe_batch(0, [pb.column("o_custkey")], want=d.OUT_PAIRS)
            g = j.gather(res, [pb.column("o_custkey")], [pb.column("o_orderkey"), pb.column("o_comment")])
            cols = g.to_arrow([pa.int64(), pa.string(), pa.int32(), pa.int64(), pa.int64(), pa.string()])
            pieces.append(pa.Table.from_arrays(cols, names=["c_custkey", "c_name", "c_mktsegment", "o_custkey", "o_orderkey", "o_comment"]))
            g.release()
            res.release()
    got = pa.concat_tables(pieces)
    want = orders.join(build, keys="o_custkey", right_keys="c_custkey", join_type="inner", coalesce_keys=False)
    want = want.select(["c_custkey", "c_name", "c_mktsegment", "o_custkey", "o_orderkey", "o_comment"])
    assert got.num_rows == want.num_rows
    assert rows_multiset(got) == rows_multiset(want)
