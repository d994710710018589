"""Output ROWS of a join over several ranks (SURVEY §8e + a6: take_multiple_record_batch, src/shared/shared.rs:83-92, when the
columns stayed on the ranks that own the rows). All ranks live on cuda:0 here -- plain device pointers stand for the cudaIpc
peer pointers of a real node (tests/test_multi_gpu.py runs the same through DistributedJoin on 2 GPUs) -- so the 1-GPU
suite covers: the exchange of (key, row id), the local joins, b200j_resolve_origin + b200j_gather_pieces(_utf8) of Int64 /
Int32 / Utf8 payload with NULLs for every join type's row shape, and composite / Utf8 join KEYS through the exchange (the
exchange carries the row hash; candidate pairs are checked against the key columns at their sources and go back through
b200j_apply_pairs). Compared, as row multisets, with the oracle's whole-join output on the concatenated tables."""
import numpy as np
import pyarrow as pa
import pytest

import oracle
from helpers import rows_multiset

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    import torch
    import datafusion_parallelism_b200 as d
    assert d.device_count() >= 1, "no CUDA device: -m gpu tests must run on the B200 box"
    return d, torch


def _tables(world, rng, nb, npr, key_space, null_keys, key_kind):
    """Per-rank build / probe tables: key column(s) + an Int64 payload with NULLs + a Utf8 payload with NULLs + an Int32 payload."""
    def keycols(n, side):
        k = rng.integers(0, key_space, n)
        if key_kind == "int64":
            cols = {"k": pa.array(k.astype(np.int64), mask=(rng.random(n) < 0.05) if null_keys else None)}
        elif key_kind == "composite":
            cols = {"k": pa.array((k // 7).astype(np.int64), mask=(rng.random(n) < 0.05) if null_keys else None),
                    "k2": pa.array((k % 7).astype(np.int32))}
        else:   # utf8 (+ an int32 second key)
            cols = {"k": pa.array([f"key-{v // 3:05d}" + "x" * int(v % 5) for v in k], mask=(rng.random(n) < 0.05) if null_keys else None),
                    "k2": pa.array((k % 3).astype(np.int32))}
        return cols
    build, probe = [], []
    for r in range(world):
        for n, out, tag in ((nb, build, "b"), (npr, probe, "p")):
            cols = keycols(n, tag)
            cols["v"] = pa.array(rng.integers(-10**12, 10**12, n).astype(np.int64), mask=rng.random(n) < 0.1)
            cols["s"] = pa.array([f"{tag}{r}-{i}" * (1 + i % 3) for i in range(n)], mask=rng.random(n) < 0.1)
            cols["w"] = pa.array(rng.integers(-1000, 1000, n).astype(np.int32))
            out.append(pa.table(cols))
    return build, probe


class _Rank:
    pass


def _share(torch, rows, tables, names):
    """SharedColumn per column name: the per-rank device buffers (plain pointers: every 'rank' is on this GPU)."""
    out = {}
    for name in names:
        devs = [rows.arrow_to_device(torch, t.column(name)) for t in tables]
        out[name] = rows.SharedColumn(devs[0]["type"], [d_["values"].data_ptr() for d_ in devs],
                                      [d_["validity"].data_ptr() if d_["validity"] is not None else 0 for d_ in devs],
                                      [d_["data"].data_ptr() if d_["data"] is not None else 0 for d_ in devs], keep=devs)
    return out


def _exchange(d, torch, joins, key_tensors, validity, world):
    """(key, row id) of every rank's rows to their destination ranks (two-pass count + scatter into plain device buffers).
    Returns per rank: received keys tensor, OriginPieces, and the (null-key row ids tensor) kept at the source."""
    from datafusion_parallelism_b200.exchange import ExchangePlan, shuffle_count, shuffle_scatter
    from datafusion_parallelism_b200.join import ColumnRef
    from datafusion_parallelism_b200 import rows
    cols = [ColumnRef(key_tensors[r].data_ptr(), None, validity[r].data_ptr() if validity[r] is not None else None, 0,
                      int(key_tensors[r].shape[0]), d.MEM_DEVICE, 8, (key_tensors[r], validity[r])) for r in range(world)]
    counts = [shuffle_count(joins[r], cols[r], world) for r in range(world)]
    matrix = np.stack([c[:world] for c in counts])
    plans = [ExchangePlan(matrix, r) for r in range(world)]
    recv_k = [torch.empty(int(plans[r].recv_rows[r]) + 8, dtype=torch.int64, device="cuda") for r in range(world)]
    recv_r = [torch.empty(int(plans[r].recv_rows[r]) + 8, dtype=torch.int32, device="cuda") for r in range(world)]
    null_r = []
    torch.cuda.synchronize()
    for r in range(world):
        n_null = int(counts[r][world])
        nk = torch.empty(n_null + 8, dtype=torch.int64, device="cuda")
        nr = torch.empty(n_null + 8, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        c = shuffle_count(joins[r], cols[r], world)
        assert np.array_equal(c, counts[r])
        shuffle_scatter(joins[r], world, [t.data_ptr() for t in recv_k] + [nk.data_ptr()], [t.data_ptr() for t in recv_r] + [nr.data_ptr()],
                        list(plans[r].dest_offset) + [0])
        joins[r].synchronize()
        null_r.append(nr[:n_null])
    origins = []
    for r in range(world):
        o = rows.OriginPieces()
        for s in range(world):
            o.add(s, int(matrix[s, r]), recv_r[r].data_ptr() + int(plans[r].src_offset[s]) * 4, keep=recv_r[r])
        origins.append(o)
    return [recv_k[r][: int(plans[r].recv_rows[r])] for r in range(world)], origins, null_r


def _hash_rows(d, torch, hasher, table, key_names):
    """b200j_hash_keys over the key columns of one rank's table -> int64 device tensor (the 8-byte 'key' the exchange carries)."""
    h = hasher.hash_keys([table.column(n) for n in key_names]).view(np.int64)
    return torch.from_numpy(h.copy()).cuda()


@pytest.mark.parametrize("join_type", ["inner", "left", "right", "full", "left_semi", "left_anti", "right_semi", "right_anti"])
@pytest.mark.parametrize("world,key_kind", [(2, "int64"), (3, "int64"), (3, "composite"), (2, "utf8")])
def test_join_rows_over_emulated_ranks(env, join_type, world, key_kind):
    d, torch = env
    from datafusion_parallelism_b200 import rows
    rng = np.random.default_rng(world * 100 + len(join_type) + len(key_kind))
    nb, npr = 3_000, 9_000
    build, probe = _tables(world, rng, nb, npr, key_space=4_000, null_keys=True, key_kind=key_kind)
    key_names = ["k"] if key_kind == "int64" else ["k", "k2"]
    pay_names = ["v", "s", "w"]
    hashed = key_kind != "int64"
    left_only, right_only = join_type in ("left_semi", "left_anti"), join_type in ("right_semi", "right_anti")
    joins = [d.HashJoin(join_type, key_widths=(8,), layout="radix", radix_bits=3) for _ in range(world)]
    hasher = d.HashJoin("inner", key_widths=tuple(d.T_UTF8 if pa.types.is_string(build[0].column(n).type) else
                                                  build[0].column(n).type.bit_width // 8 for n in key_names)) if hashed else None
    try:
        def side_keys(tables):
            if hashed:       # NULL-key rows travel like any other row: the key check drops their candidate pairs
                return [_hash_rows(d, torch, hasher, t, key_names) for t in tables], [None] * world
            devs = [rows.arrow_to_device(torch, t.column("k")) for t in tables]
            return [x["values"].view(torch.int64) for x in devs], [x["validity"] for x in devs]

        bkeys, bvalid = side_keys(build)
        recv_bk, origin_b, null_brid = _exchange(d, torch, joins, bkeys, bvalid, world)
        for r in range(world):
            joins[r].build_push(0, [recv_bk[r]])
            joins[r].build_finish(0)
        pkeys, pvalid = side_keys(probe)
        recv_pk, origin_p, null_rid = _exchange(d, torch, joins, pkeys, pvalid, world)
        bcols = _share(torch, rows, build, key_names + pay_names)
        pcols = _share(torch, rows, probe, key_names + pay_names)
        out_types = {n: build[0].column(n).type for n in key_names + pay_names}
        pieces = []

        def emit(j, gb, gp):
            cols, names = [], []
            if not right_only:
                for n in key_names + pay_names:
                    cols.append(gb.gather(bcols[n]).to_arrow(out_types[n]))
                    names.append(f"l.{n}")
            if not left_only:
                for n in key_names + pay_names:
                    cols.append(gp.gather(pcols[n]).to_arrow(out_types[n]))
                    names.append(f"r.{n}")
            pieces.append(pa.Table.from_arrays(cols, names=names))

        for r in range(world):
            j = joins[r]
            n_recv = int(recv_pk[r].shape[0])
            if not hashed:
                res = j.probe_batch(0, [recv_pk[r]], want=d.OUT_PAIRS)
            else:
                cand = j.probe_batch(0, [recv_pk[r]], want=d.OUT_PAIRS | d.OUT_MATCHES_ONLY)
                cb, cp = cand.device_arrays()
                if cand.n:
                    cb_t, cp_t = torch.as_tensor(cb, device="cuda"), torch.as_tensor(cp, device="cuda")
                    gb = rows.RowGatherer(j, torch, cb_t, cand.n, 8, origin_b[r])
                    gp = rows.RowGatherer(j, torch, cp_t, cand.n, 4, origin_p[r])
                    keep = rows.equal_key_rows(j, torch, gb, gp, [bcols[n] for n in key_names], [pcols[n] for n in key_names])
                    kb, kp = cb_t[keep].contiguous(), cp_t[keep].contiguous()
                else:
                    kb = torch.empty(0, dtype=torch.int64, device="cuda")
                    kp = torch.empty(0, dtype=torch.int32, device="cuda")
                torch.cuda.synchronize()
                res = j.apply_pairs_device(0, n_recv, kb.data_ptr(), kp.data_ptr(), int(kb.shape[0]), want=d.OUT_PAIRS)
                cand.release()
            if res.n:
                b_, p_ = res.device_arrays()
                gb = rows.RowGatherer(j, torch, torch.as_tensor(b_, device="cuda"), res.n, res.build_idx_width, origin_b[r]) \
                    if b_ is not None else rows.RowGatherer.all_null(j, torch, res.n)
                gp = rows.RowGatherer(j, torch, torch.as_tensor(p_, device="cuda"), res.n, 4, origin_p[r]) \
                    if p_ is not None else rows.RowGatherer.all_null(j, torch, res.n)
                emit(j, gb, gp)
            res.release()
            fin = j.probe_finish(0, want=d.OUT_PAIRS)
            if fin.n:
                b_, _ = fin.device_arrays()
                emit(j, rows.RowGatherer(j, torch, torch.as_tensor(b_, device="cuda"), fin.n, fin.build_idx_width, origin_b[r]),
                     rows.RowGatherer.all_null(j, torch, fin.n))
            fin.release()
            # NULL-key probe rows never left their source rank: Right / Full / RightAnti emit them there, build side NULL
            if not hashed and join_type in ("right", "full", "right_anti") and int(null_rid[r].shape[0]):
                m = int(null_rid[r].shape[0])
                o = rows.OriginPieces()
                o.add(r, m, null_rid[r].data_ptr())
                ident = torch.arange(m, dtype=torch.int32, device="cuda")
                emit(j, rows.RowGatherer.all_null(j, torch, m), rows.RowGatherer(j, torch, ident, m, 4, o))
            # ... and so do NULL-key build rows: Left / Full / LeftAnti emit them at their source, probe side NULL
            if not hashed and join_type in ("left", "full", "left_anti") and int(null_brid[r].shape[0]):
                m = int(null_brid[r].shape[0])
                o = rows.OriginPieces()
                o.add(r, m, null_brid[r].data_ptr())
                ident = torch.arange(m, dtype=torch.int32, device="cuda")
                emit(j, rows.RowGatherer(j, torch, ident, m, 4, o), rows.RowGatherer.all_null(j, torch, m))
    finally:
        for j in joins:
            j.close()
        if hasher is not None:
            hasher.close()
    got = pa.concat_tables(pieces) if pieces else None
    want = oracle.join_tables(pa.concat_tables(build), [pa.concat_tables(probe)], [(n, n) for n in key_names], join_type)
    if got is None:
        assert want.num_rows == 0
        return
    assert got.num_rows == want.num_rows
    assert rows_multiset(got.select(want.column_names)) == rows_multiset(want)


def test_resolve_origin_and_gather_edge_cases(env):
    """NULL / out-of-range indices, identity pieces with a base, 1/2/16-byte values, empty inputs."""
    d, torch = env
    from datafusion_parallelism_b200 import rows
    with d.HashJoin("inner", key_widths=(8,)) as j:
        rid0 = torch.tensor([5, 4, 3], dtype=torch.int32, device="cuda")
        o = rows.OriginPieces()
        o.add(1, 3, rid0.data_ptr())           # piece 0: source 1, explicit row ids
        o.add(0, 0, 0)                         # an empty piece
        o.add(2, 4, 0, rid_base=10)            # piece 2: identity + base -> rows 10..13 of source 2
        idx = torch.tensor([0, 2, 3, 6, 7, -1, 1], dtype=torch.int64, device="cuda")
        g = rows.RowGatherer(j, torch, idx, 7, 8, o)
        assert g.piece.cpu().tolist()[:7] == [0, 0, 2, 2, 0, 0, 0]
        assert (g.row.cpu().numpy().view(np.uint32)[:7]).tolist() == [5, 3, 10, 13, 0xFFFFFFFF, 0xFFFFFFFF, 4]
        for width, dt in ((1, torch.uint8), (2, torch.int16), (16, torch.int64)):
            srcs = [(torch.arange(40, dtype=torch.int64, device="cuda") * (s + 1)).to(dt) if width != 16 else
                    (torch.arange(80, dtype=torch.int64, device="cuda") * (s + 1)).reshape(40, 2).contiguous() for s in range(3)]
            col = rows.SharedColumn(width, [t.data_ptr() for t in srcs])
            out = g.gather(col)
            valid = out.validity_numpy()
            assert valid.tolist() == [True, True, True, True, False, False, True]
            exp_rank_row = [(1, 5), (1, 3), (2, 10), (2, 13), None, None, (1, 4)]
            vals = out.values.cpu()
            for i, rr in enumerate(exp_rank_row):
                if rr is None:
                    continue
                assert torch.equal(vals[i], srcs[rr[0]][rr[1]].cpu())
        # u32 indices, empty input
        g0 = rows.RowGatherer(j, torch, torch.empty(0, dtype=torch.int32, device="cuda"), 0, 4, o)
        assert g0.gather(rows.SharedColumn(8, [0, 0, 0])).n == 0
        # more pieces than the table holds is refused
        big = rows.OriginPieces()
        for s in range(65):
            big.add(0, 1, 0)
        with pytest.raises(ValueError):
            rows.RowGatherer(j, torch, idx, 7, 8, big)
