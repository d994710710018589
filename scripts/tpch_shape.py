"""C5 (BASELINE.json configs[4]): the join trees of TPC-H Q3 / Q5 / Q9 run through the GPU join operator (HashJoin:
b200j_build_push / build_finish / probe_batch / gather), on TPC-H-SHAPED data generated on the device.

NOT dbgen-identical: the reference generates SF100 with `tpchgen-cli` (tpc/scripts/generate.sh:82-103, cargo-installed) or a
docker dbgen image; neither exists here (no cargo, no docker, no network). This generator keeps what the joins see — table
cardinalities, the sparse o_orderkey space, 1..7 lineitems per order, customers without orders, the partsupp
(partkey, suppkey) formula lineitem draws its supplier from, uniform nation keys, the date ranges and the filter
selectivities of the three queries — and nothing else (no strings, no prices). Filters are torch masks in front of the joins
(DataFusion pushes them below the joins as well); aggregation is left out: the output of the last join is counted and
checksummed. `--verify` re-runs every join with pyarrow Acero on the host (small scale factors only).

Queries: tpc/datafusion-benchmarks/tpch/queries/q3.sql, q5.sql, q9.sql.
"""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import datafusion_parallelism_b200 as d
from datafusion_parallelism_b200.exchange import _RawCudaArray

ap = argparse.ArgumentParser()
ap.add_argument("--sf", type=float, default=100.0)
ap.add_argument("--queries", default="q3,q5,q9")
ap.add_argument("--verify", action="store_true")
ap.add_argument("--chunk", type=int, default=1 << 28)
ap.add_argument("--hint", action="store_true", help="probe with B200J_OUT_GATHER_HINT (measured neutral on these join trees: most joins are filtered, their results compacted)")
a = ap.parse_args()
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
I64 = torch.int64
DAYS = 2406                        # 1992-01-01 .. 1998-08-02 (o_orderdate range)
D_1995_03_15 = 1169
D_1994_01_01, D_1995_01_01 = 731, 1096


def gen(sf):
    g = torch.Generator(device=dev)
    g.manual_seed(42)
    n_cust, n_ord, n_part, n_supp = int(150_000 * sf), int(1_500_000 * sf), int(200_000 * sf), max(int(10_000 * sf), 4)
    t = {}
    t["c_custkey"] = torch.arange(1, n_cust + 1, device=dev, dtype=I64)
    t["c_nationkey"] = torch.randint(0, 25, (n_cust,), device=dev, generator=g, dtype=I64)
    t["c_mktsegment"] = torch.randint(0, 5, (n_cust,), device=dev, generator=g, dtype=I64)
    i = torch.arange(n_ord, device=dev, dtype=I64)
    t["o_orderkey"] = (i // 8) * 32 + (i % 8) + 1                                   # dbgen: 8 keys used out of every 32
    ck = torch.randint(1, n_cust + 1, (n_ord,), device=dev, generator=g, dtype=I64)
    t["o_custkey"] = torch.where(ck % 3 == 0, (ck % n_cust) + 1, ck)                # a third of the customers have no orders
    t["o_custkey"] = torch.where(t["o_custkey"] % 3 == 0, torch.clamp(t["o_custkey"] - 1, min=1), t["o_custkey"])
    t["o_orderdate"] = torch.randint(0, DAYS, (n_ord,), device=dev, generator=g, dtype=I64)
    nl = torch.randint(1, 8, (n_ord,), device=dev, generator=g, dtype=I64)          # 1..7 lineitems per order
    t["l_orderkey"] = torch.repeat_interleave(t["o_orderkey"], nl)
    l_odate = torch.repeat_interleave(t["o_orderdate"], nl)
    n_li = int(t["l_orderkey"].numel())
    t["l_shipdate"] = l_odate + torch.randint(1, 122, (n_li,), device=dev, generator=g, dtype=I64)
    del l_odate, nl, i, ck
    t["l_partkey"] = torch.randint(1, n_part + 1, (n_li,), device=dev, generator=g, dtype=I64)
    jj = torch.randint(0, 4, (n_li,), device=dev, generator=g, dtype=I64)
    S = n_supp
    t["l_suppkey"] = (t["l_partkey"] + jj * (S // 4 + (t["l_partkey"] - 1) // S)) % S + 1      # dbgen's partsupp supplier formula
    del jj
    pp = torch.arange(1, n_part + 1, device=dev, dtype=I64).repeat_interleave(4)
    pj = torch.arange(0, 4, device=dev, dtype=I64).repeat(n_part)
    t["ps_partkey"] = pp
    t["ps_suppkey"] = (pp + pj * (S // 4 + (pp - 1) // S)) % S + 1
    t["ps_supplycost"] = torch.randint(100, 100_000, (4 * n_part,), device=dev, generator=g, dtype=I64)
    del pj
    t["p_partkey"] = torch.arange(1, n_part + 1, device=dev, dtype=I64)
    t["p_green"] = torch.rand(n_part, device=dev, generator=g) < 5.0 / 92.0        # p_name like '%green%': 5 words out of 92 colours
    t["s_suppkey"] = torch.arange(1, n_supp + 1, device=dev, dtype=I64)
    t["s_nationkey"] = torch.randint(0, 25, (n_supp,), device=dev, generator=g, dtype=I64)
    t["n_nationkey"] = torch.arange(0, 25, device=dev, dtype=I64)
    t["n_regionkey"] = torch.tensor([0, 1, 1, 1, 4, 0, 3, 3, 2, 2, 4, 4, 2, 4, 0, 0, 0, 1, 2, 3, 4, 2, 3, 3, 1], device=dev, dtype=I64)
    torch.cuda.synchronize()
    return t


LOG = []


def join(name, bkeys, bpay, pkeys, ppay):
    """Inner join through the operator: build (keys, payload), probe (keys, payload) in chunks, gather on the device.
    Returns the output columns [build keys..., build payload..., probe keys..., probe payload...] as torch tensors."""
    nk = len(bkeys)
    t0 = time.perf_counter()
    outs = None
    with d.HashJoin("inner", key_widths=(8,) * nk, build_payload_types=(8,) * len(bpay), probe_payload_types=(8,) * len(ppay),
                    timing=True, expected_build_rows=int(bkeys[0].numel())) as j:
        j.build_push(0, bkeys, bpay)
        j.build_finish(0)
        n_probe = int(pkeys[0].numel())
        n_cols = 2 * nk + len(bpay) + len(ppay)
        pieces = [[] for _ in range(n_cols)]
        for lo in range(0, max(n_probe, 1), a.chunk):
            pk = [k[lo:lo + a.chunk] for k in pkeys]
            pp = [p[lo:lo + a.chunk] for p in ppay]
            res = j.probe_batch(0, pk, want=d.OUT_PAIRS | (d.OUT_GATHER_HINT if a.hint else 0))
            gth = j.gather(res, pk, pp, want=0)                                      # device result
            n = gth.n_rows
            for c in range(n_cols):
                ptr = gth.column_buffers(c)["values"]
                pieces[c].append(torch.as_tensor(_RawCudaArray(ptr, n, "<i8"), device=dev).clone() if n else torch.empty(0, dtype=I64, device=dev))
            gth.release()
            res.release()
        st = j.stats()
        outs = [torch.cat(p) for p in pieces]
    wall = (time.perf_counter() - t0) * 1e3
    ms = {k: round(st[k], 3) for k in ("ms_build_copy", "ms_build_partition", "ms_build_insert", "ms_probe_partition", "ms_probe", "ms_gather", "ms_slot_order")}
    LOG.append({"join": name, "keys": nk, "build_rows": int(bkeys[0].numel()), "probe_rows": int(pkeys[0].numel()), "out_rows": int(outs[0].numel()),
                "device_ms": round(sum(ms.values()), 3), "wall_ms": round(wall, 2), "layout": st["layout"], **ms})
    if a.verify:
        import pyarrow as pa
        bt = pa.table({f"k{i}": bkeys[i].cpu().numpy() for i in range(nk)})
        pt = pa.table({f"k{i}": pkeys[i].cpu().numpy() for i in range(nk)})
        ref = pt.join(bt.group_by([f"k{i}" for i in range(nk)]).aggregate([([], "count_all")]), keys=[f"k{i}" for i in range(nk)], join_type="inner")
        want = int(np.sum(ref.column("count_all").to_numpy()))
        assert want == int(outs[0].numel()), (name, want, int(outs[0].numel()))
        for i in range(nk):                                                         # key equality on every output row
            assert bool((outs[i] == outs[nk + len(bpay) + i]).all()), name
    return outs


def q3(t):
    c = t["c_custkey"][t["c_mktsegment"] == 0]
    om = t["o_orderdate"] < D_1995_03_15
    lm = t["l_shipdate"] > D_1995_03_15
    # customer x orders on custkey -> (o_orderkey, o_orderdate)
    o = join("q3 customer[BUILDING] x orders[date<1995-03-15]", [c], [], [t["o_custkey"][om]], [t["o_orderkey"][om], t["o_orderdate"][om]])
    okey, odate = o[2], o[3]
    # x lineitem on orderkey
    r = join("q3 .. x lineitem[shipdate>1995-03-15]", [okey], [odate], [t["l_orderkey"][lm]], [t["l_partkey"][lm]])
    return int(r[0].numel()), int((r[0].sum() + r[1].sum()).item())


def q5(t):
    asia = t["n_nationkey"][t["n_regionkey"] == 2]
    c = join("q5 nation[ASIA] x customer", [asia], [], [t["c_nationkey"]], [t["c_custkey"]])
    c_nat, c_key = c[1], c[2]
    om = (t["o_orderdate"] >= D_1994_01_01) & (t["o_orderdate"] < D_1995_01_01)
    o = join("q5 .. x orders[1994]", [c_key], [c_nat], [t["o_custkey"][om]], [t["o_orderkey"][om]])
    o_nat, o_key = o[1], o[3]
    l = join("q5 .. x lineitem", [o_key], [o_nat], [t["l_orderkey"]], [t["l_suppkey"]])
    l_nat, l_supp = l[1], l[3]
    # supplier on (suppkey, nationkey): c_nationkey = s_nationkey is part of the join key
    r = join("q5 supplier x .. on (suppkey, nationkey)", [t["s_suppkey"], t["s_nationkey"]], [], [l_supp, l_nat], [])
    return int(r[0].numel()), int((r[0].sum() + r[1].sum()).item())


def q9(t):
    green = t["p_partkey"][t["p_green"]]
    l = join("q9 part[green] x lineitem", [green], [], [t["l_partkey"]], [t["l_suppkey"], t["l_orderkey"]])
    l_part, l_supp, l_ord = l[1], l[2], l[3]
    s = join("q9 supplier x ..", [t["s_suppkey"]], [t["s_nationkey"]], [l_supp], [l_part, l_ord])
    s_nat, l_supp, l_part, l_ord = s[1], s[2], s[3], s[4]
    ps = join("q9 partsupp x .. on (partkey, suppkey)", [t["ps_partkey"], t["ps_suppkey"]], [t["ps_supplycost"]], [l_part, l_supp], [l_ord, s_nat])
    cost, l_ord, s_nat = ps[2], ps[5], ps[6]
    o = join("q9 orders x ..", [t["o_orderkey"]], [t["o_orderdate"]], [l_ord], [cost, s_nat])
    r = join("q9 nation x ..", [t["n_nationkey"]], [], [o[4]], [o[1], o[3]])
    return int(r[0].numel()), int((r[2].sum() + r[3].sum()).item())


t0 = time.perf_counter()
T = gen(a.sf)
gen_s = time.perf_counter() - t0
sizes = {k: int(v.numel()) for k, v in T.items() if k in ("c_custkey", "o_orderkey", "l_orderkey", "ps_partkey", "p_partkey", "s_suppkey")}
for q in a.queries.split(","):
    LOG.clear()
    for rep in range(2):                                # second run: allocations warm
        LOG.clear()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        rows, checksum = {"q3": q3, "q5": q5, "q9": q9}[q](T)
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) * 1e3
    print(json.dumps({"query": q, "sf": a.sf, "data": "TPC-H-shaped (not dbgen-identical), generated on the device", "tables": sizes,
                      "result_rows": rows, "checksum": checksum, "join_device_ms": round(sum(x["device_ms"] for x in LOG), 2),
                      "wall_ms_incl_filters_and_column_copies": round(wall, 1), "verified_vs_acero": bool(a.verify), "joins": list(LOG)}), flush=True)
