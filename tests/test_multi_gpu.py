"""Real multi-GPU parity (runs when the box has >= 2 GPUs; the 1-GPU CI skips it): one process per GPU, both sides
exchanged over NVLink in every exchange mode, results mapped back to global row ids through the origin mapping and compared
with the single-table oracle join as order-normalised multisets. The N > 1 host logic is covered on CPU by
tests/test_exchange.py (gloo)."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, mode, join_type, unique, out_dir):
    import torch
    import torch.distributed as dist
    import datafusion_parallelism_b200 as d
    from datafusion_parallelism_b200.exchange import DistributedJoin
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    rng = np.random.default_rng(100 + rank)
    nb, npr = 40_000, 150_000
    if unique:
        bk = (rng.permutation(nb * 2)[:nb].astype(np.int64)) * world + rank            # unique over all ranks
    else:
        bk = rng.integers(0, nb // 2, nb).astype(np.int64)
    pk = rng.integers(-100, nb * 2 * world + 100, npr).astype(np.int64)
    if mode == "pull-wide":
        pk[7] = 2**40 + rank                                                            # forces the 8-byte fall-back on every rank
    stream = torch.cuda.Stream(device=dev)
    j = DistributedJoin(join_type, key_width=8, stream=stream, mode=mode.split("-")[0], layout="radix", radix_bits=4)
    out = {}
    for rep in range(2):                                                                # second pass re-uses every buffer
        j.reset()
        j.build_push(0, [torch.from_numpy(bk).to(dev)])
        j.build_finish(0)
        r = j.probe_batch(0, [torch.from_numpy(pk).to(dev)], want=d.OUT_PAIRS)
        b_, p_ = r.device_arrays()
        lb = torch.as_tensor(b_, device=dev).cpu().numpy().view(np.uint64) if r.n and b_ is not None else np.zeros(0, np.uint64)
        lp = torch.as_tensor(p_, device=dev).cpu().numpy().view(np.uint32) if r.n and p_ is not None else np.zeros(0, np.uint32)
        r.release()
        NULLB, NULLP = np.uint64(0xFFFFFFFFFFFFFFFF), np.uint32(0xFFFFFFFF)
        gb = np.full(len(lb), NULLB, dtype=np.uint64)
        m = lb != NULLB
        if m.any():
            src, row = j.origin_build(lb[m].astype(np.int64))
            gb[m] = src.astype(np.uint64) * nb + row
        psrc, prow = j.origin_probe(lp.astype(np.int64))
        gp = psrc.astype(np.uint64) * npr + prow
        fin = j.probe_finish(0, want=d.OUT_PAIRS | d.OUT_HOST)
        fb, _ = fin.numpy()
        fin.release()
        fsrc, frow = j.origin_build(fb.astype(np.int64)) if len(fb) else (np.zeros(0, np.int64), np.zeros(0, np.uint32))
        out = {"gb": gb, "gp": gp, "fin": fsrc.astype(np.uint64) * nb + frow, "bk": bk, "pk": pk}
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), **out)
    dist.barrier()
    j.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["pull", "pull-wide", "p2p", "ce"])
@pytest.mark.parametrize("join_type,unique", [("inner", True), ("inner", False), ("full", True)])
def test_two_gpu_join_matches_oracle(mode, join_type, unique, tmp_path):
    import torch
    import torch.multiprocessing as mp
    import oracle
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), mode, join_type, unique, str(tmp_path)), nprocs=world, join=True)
    parts = [np.load(tmp_path / f"rank{r}.npz") for r in range(world)]
    bk = np.concatenate([p["bk"] for p in parts])
    pk = np.concatenate([p["pk"] for p in parts])
    oj = oracle.OracleJoin(join_type, [bk])
    wb, wp = oj.probe_batch([pk], len(pk))
    wfin, _ = oj.finish()
    gb = np.concatenate([p["gb"] for p in parts])
    gp = np.concatenate([p["gp"] for p in parts])
    # the oracle returns the NULL-key probe rows of a Full join too; the exchange keeps them on their source rank (never sent)
    o1, o2 = np.lexsort((gb, gp)), np.lexsort((wb, wp.astype(np.uint64)))
    assert gb.shape == wb.shape
    assert np.array_equal(gb[o1], wb[o2]) and np.array_equal(gp[o1], wp.astype(np.uint64)[o2])
    gfin = np.sort(np.concatenate([p["fin"] for p in parts]))
    assert np.array_equal(gfin, np.sort(wfin))


def _rows_tables(rank, keyed):
    """This rank's build / probe tables: key column(s), an Int64 payload with NULLs, a Utf8 payload with NULLs."""
    import pyarrow as pa
    rng = np.random.default_rng(500 + rank)
    nb, npr = 20_000, 60_000
    out = []
    for n, tag in ((nb, "b"), (npr, "p")):
        k = rng.integers(0, 30_000, n)
        if keyed == "int64":
            cols = {"k": pa.array(k.astype(np.int64))}
        else:       # a Utf8 key + an Int32 key, NULLs in the first
            cols = {"k": pa.array([f"key-{v // 3:06d}" for v in k], mask=rng.random(n) < 0.03), "k2": pa.array((k % 3).astype(np.int32))}
        cols["v"] = pa.array(rng.integers(-10**12, 10**12, n).astype(np.int64), mask=rng.random(n) < 0.1)
        cols["s"] = pa.array([f"{tag}{rank}-{i}" * (1 + i % 3) for i in range(n)], mask=rng.random(n) < 0.1)
        out.append(pa.table(cols))
    return out


def _rows_worker(rank, world, port, mode, join_type, keyed, out_dir):
    import pyarrow as pa
    import pyarrow.feather as feather
    import torch
    import torch.distributed as dist
    import datafusion_parallelism_b200 as d
    from datafusion_parallelism_b200.exchange import DistributedJoin, DistributedKeyJoin
    from datafusion_parallelism_b200.rows import RowGatherer
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    build, probe = _rows_tables(rank, keyed)
    key_names = ["k"] if keyed == "int64" else ["k", "k2"]
    names = build.column_names
    stream = torch.cuda.Stream(device=dev)
    if keyed == "int64":
        kj = None
        j = DistributedJoin(join_type, key_width=8, stream=stream, mode=mode, layout="radix", radix_bits=4, track_origin=True)
        half = build.num_rows // 2              # two build batches: the row ids count this rank's build rows over both
        for lo, hi in ((0, half), (half, build.num_rows)):
            j.build_push(0, [torch.from_numpy(np.asarray(build.column("k"))[lo:hi].copy()).to(dev)])
        j.build_finish(0)
        res = j.probe_batch(0, [torch.from_numpy(np.asarray(probe.column("k")).copy()).to(dev)], want=d.OUT_PAIRS)
    else:
        kj = DistributedKeyJoin(join_type, key_types=(d.T_UTF8, 4), stream=stream, mode=mode, layout="radix", radix_bits=4)
        j = kj.dj
        kj.build_push(0, [build.column(n) for n in key_names])
        kj.build_finish(0)
        res = kj.probe_batch(0, [probe.column(n) for n in key_names], want=d.OUT_PAIRS)
    bcols = j.share_columns([build.column(n) for n in names])
    pcols = j.share_columns([probe.column(n) for n in names])
    pieces = []

    def emit(gb, gp):
        cols = [gb.gather(c).to_arrow(build.column(n).type) for n, c in zip(names, bcols)]
        cols += [gp.gather(c).to_arrow(probe.column(n).type) for n, c in zip(names, pcols)]
        pieces.append(pa.Table.from_arrays(cols, names=[f"l.{n}" for n in names] + [f"r.{n}" for n in names]))

    if res.n:
        b_, p_ = res.device_arrays()
        emit(j.build_rows(torch.as_tensor(b_, device=dev), res.n, res.build_idx_width), j.probe_rows(torch.as_tensor(p_, device=dev), res.n))
    res.release()
    fin = j.probe_finish(0, want=d.OUT_PAIRS)
    if fin.n:
        b_, _ = fin.device_arrays()
        emit(j.build_rows(torch.as_tensor(b_, device=dev), fin.n, fin.build_idx_width), RowGatherer.all_null(j.join, torch, fin.n))
    fin.release()
    out = pa.concat_tables(pieces) if pieces else pa.table({})
    feather.write_feather(out, os.path.join(out_dir, f"out{rank}.feather"))
    feather.write_feather(build, os.path.join(out_dir, f"build{rank}.feather"))
    feather.write_feather(probe, os.path.join(out_dir, f"probe{rank}.feather"))
    dist.barrier()                  # nobody unmaps columns a peer is still reading
    j.unshare_columns(bcols)
    j.unshare_columns(pcols)
    (kj or j).close()
    dist.destroy_process_group()


@pytest.mark.parametrize("mode,join_type,keyed", [("pull", "inner", "int64"), ("pull", "full", "int64"), ("p2p", "left", "int64"),
                                                  ("ce", "full", "int64"), ("pull", "inner", "utf8"), ("pull", "full", "utf8"),
                                                  ("p2p", "right", "utf8")])
def test_two_gpu_join_rows_match_oracle(mode, join_type, keyed, tmp_path):
    """Full output ROWS (Int64 + Utf8 payload with NULLs, gathered from the ranks that own them over NVLink) of a join over 2
    GPUs against the oracle's whole-join output; `utf8`: a (Utf8, Int32) composite key through the exchange."""
    import pyarrow as pa
    import pyarrow.feather as feather
    import torch
    import torch.multiprocessing as mp
    import oracle
    from helpers import rows_multiset
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    world = 2
    mp.spawn(_rows_worker, args=(world, _free_port(), mode, join_type, keyed, str(tmp_path)), nprocs=world, join=True)
    build = pa.concat_tables([feather.read_table(tmp_path / f"build{r}.feather") for r in range(world)])
    probe = pa.concat_tables([feather.read_table(tmp_path / f"probe{r}.feather") for r in range(world)])
    outs = [feather.read_table(tmp_path / f"out{r}.feather") for r in range(world)]
    got = pa.concat_tables([t for t in outs if t.num_columns])
    key_names = ["k"] if keyed == "int64" else ["k", "k2"]
    want = oracle.join_tables(build, [probe], [(n, n) for n in key_names], join_type)
    assert got.num_rows == want.num_rows
    assert rows_multiset(got.select(want.column_names)) == rows_multiset(want)
