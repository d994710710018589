"""Multi-GPU exchange: host-side planning on CPU (gloo, world_size 2 and 3) and the partition/scatter kernels on one GPU
with the ranks emulated in one process (B200_PROFILING.md: with fewer GPUs than ranks, emulate the ranks)."""
import os
import socket

import numpy as np
import pytest

import oracle


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _plan_worker(rank, world, port, seed, out_dir):
    import torch
    import torch.distributed as dist
    # importing the package needs the built library (CPU box: built by build()); the plan itself is pure host logic
    from datafusion_parallelism_b200.exchange import ExchangePlan
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(seed + rank)
    mine = torch.from_numpy(rng.integers(0, 1000, world).astype(np.int64))
    if rank == 1:
        mine[0] = 0                       # an empty segment
    allc = [torch.zeros(world, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(allc, mine)
    matrix = torch.stack(allc).numpy()
    plan = ExchangePlan(matrix, rank)
    gathered = [None] * world
    dist.all_gather_object(gathered, {"dest_offset": plan.dest_offset.tolist(), "src_offset": plan.src_offset.tolist(),
                                      "recv": int(plan.recv_rows[rank]), "send_split": plan.send_split, "recv_split": plan.recv_split})
    if rank == 0:
        np.save(os.path.join(out_dir, "matrix.npy"), matrix)
        import json
        with open(os.path.join(out_dir, "plans.json"), "w") as f:
            json.dump(gathered, f)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_exchange_plan_is_consistent_across_ranks(world, tmp_path):
    import json
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_plan_worker, args=(world, port, 11, str(tmp_path)), nprocs=world, join=True)
    matrix = np.load(tmp_path / "matrix.npy")
    plans = json.load(open(tmp_path / "plans.json"))
    for d in range(world):
        # the segments the sources write into destination d tile [0, recv_rows[d]) exactly, in rank order
        segs = sorted((plans[s]["dest_offset"][d], plans[s]["dest_offset"][d] + int(matrix[s, d])) for s in range(world))
        pos = 0
        for lo, hi in segs:
            assert lo == pos
            pos = hi
        assert pos == plans[d]["recv"] == int(matrix[:, d].sum())
        # and the destination's view of where each source starts agrees
        assert plans[d]["src_offset"][:-1] == [plans[s]["dest_offset"][d] for s in range(world)]
        assert plans[d]["recv_split"] == [int(matrix[s, d]) for s in range(world)]
        assert plans[d]["send_split"] == [int(matrix[d, s]) for s in range(world)]


@pytest.mark.gpu
@pytest.mark.parametrize("world,dtype", [(2, np.int64), (3, np.int64), (4, np.int32)])
def test_emulated_ranks_join_matches_oracle(world, dtype):
    """All ranks live on cuda:0: every rank's scatter writes into every rank's receive buffer (plain device pointers here,
    cudaIpc peer pointers on a real node), then every rank joins what it received. The union of the per-rank results,
    mapped back to global row ids, must equal the single-table oracle join."""
    import torch
    import datafusion_parallelism_b200 as d
    from datafusion_parallelism_b200.exchange import ExchangePlan, shuffle_count, shuffle_scatter, _dev_column
    rng = np.random.default_rng(5)
    nb, npr = 30_000, 100_000
    kw = np.dtype(dtype).itemsize
    tdt = torch.int64 if kw == 8 else torch.int32
    bkeys = [rng.integers(0, 50_000, nb).astype(dtype) for _ in range(world)]        # duplicates across and inside ranks
    pkeys = [rng.integers(-1000, 60_000, npr).astype(dtype) for _ in range(world)]
    joins = [d.HashJoin("inner", key_widths=(kw,), layout="radix", radix_bits=3) for _ in range(world)]
    try:
        def exchange(keys_per_rank):
            dev = [torch.from_numpy(k).cuda() for k in keys_per_rank]
            counts = [shuffle_count(joins[r], dev[r], world) for r in range(world)]
            matrix = np.stack([c[:world] for c in counts])
            plans = [ExchangePlan(matrix, r) for r in range(world)]
            recv_k = [torch.empty(int(plans[r].recv_rows[r]) + 8, dtype=tdt, device="cuda") for r in range(world)]
            recv_r = [torch.empty(int(plans[r].recv_rows[r]) + 8, dtype=torch.int32, device="cuda") for r in range(world)]
            # counts were taken rank by rank, so re-count right before each rank's scatter (count -> scatter must be adjacent)
            for r in range(world):
                c = shuffle_count(joins[r], dev[r], world)
                assert np.array_equal(c, counts[r])
                shuffle_scatter(joins[r], world, [t.data_ptr() for t in recv_k] + [0], [t.data_ptr() for t in recv_r] + [0],
                                list(plans[r].dest_offset) + [0])
                joins[r].synchronize()
            return plans, recv_k, recv_r

        bplans, bk, br = exchange(bkeys)
        for r in range(world):
            n = int(bplans[r].recv_rows[r])
            joins[r].build_push(0, [_dev_column(bk[r].data_ptr(), n, kw, keep=bk[r])])
            joins[r].build_finish(0)
        pplans, pk, pr = exchange(pkeys)
        got_b, got_p = [], []
        for r in range(world):
            n = int(pplans[r].recv_rows[r])
            res = joins[r].probe_batch(0, [_dev_column(pk[r].data_ptr(), n, kw, keep=pk[r])])
            lb, lp = res.numpy()
            res.release()
            # local -> global ids: (source rank, row id at the source)
            brid, prid = br[r].cpu().numpy().view(np.uint32), pr[r].cpu().numpy().view(np.uint32)
            bsrc, _ = bplans[r].origin_of(lb.astype(np.int64))
            psrc, _ = pplans[r].origin_of(lp.astype(np.int64))
            got_b.append(bsrc.astype(np.uint64) * nb + brid[lb.astype(np.int64)])
            got_p.append(psrc.astype(np.uint64) * npr + prid[lp.astype(np.int64)])
        got_b, got_p = np.concatenate(got_b), np.concatenate(got_p)
    finally:
        for j in joins:
            j.close()
    oj = oracle.OracleJoin("inner", [np.concatenate(bkeys)])
    ob, op = oj.probe_batch([np.concatenate(pkeys)], world * npr)
    assert got_b.shape == ob.shape
    o1 = np.lexsort((got_b, got_p))
    o2 = np.lexsort((ob, op.astype(np.uint64)))
    assert np.array_equal(got_b[o1], ob[o2]) and np.array_equal(got_p[o1], op.astype(np.uint64)[o2])


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [np.int64, np.int32])
def test_hot_keys_stay_local(dtype):
    """Heavy-hitter routing of the exchange kernels (b200j_shuffle_set_hot): rows whose key is in the hot set are counted in,
    and scattered to, the extra rank-local destination; every other row goes where it went without a hot set."""
    import torch
    import datafusion_parallelism_b200 as d
    from datafusion_parallelism_b200.exchange import shuffle_count, shuffle_scatter, shuffle_set_hot
    rng = np.random.default_rng(9)
    kw, world, n = np.dtype(dtype).itemsize, 3, 200_000
    tdt = torch.int64 if kw == 8 else torch.int32
    keys = rng.integers(1, 40_000, n).astype(dtype)
    hot_vals = np.array([7, 31_337, 0xF0000001 if kw == 4 else 123_456_789_012], dtype=np.uint64)
    keys[rng.random(n) < 0.2] = dtype(7)
    keys[rng.random(n) < 0.05] = np.array([hot_vals[2]], dtype=np.uint64).astype(dtype)[0]
    is_hot = np.isin(keys.astype(np.uint64) & (0xFFFFFFFF if kw == 4 else 0xFFFFFFFFFFFFFFFF), hot_vals)
    with d.HashJoin("inner", key_widths=(kw,)) as j:
        dev = torch.from_numpy(keys).cuda()
        plain = shuffle_count(j, dev, world)
        # discard the armed count (a count must be followed by its scatter): scatter into scratch buffers
        sk = [torch.empty(n, dtype=tdt, device="cuda") for _ in range(world + 2)]
        sr = [torch.empty(n, dtype=torch.int32, device="cuda") for _ in range(world + 2)]
        shuffle_scatter(j, world, [t.data_ptr() for t in sk[: world + 1]], [t.data_ptr() for t in sr[: world + 1]], [0] * (world + 1))
        j.synchronize()
        dest_plain = np.full(n, -1)
        for r in range(world):
            dest_plain[sr[r][: int(plain[r])].cpu().numpy().view(np.uint32)] = r
        assert (dest_plain >= 0).all()

        shuffle_set_hot(j, hot_vals)
        c = shuffle_count(j, dev, world, hot=True)
        assert int(c[world + 1]) == int(is_hot.sum()) and int(c[world]) == 0
        for r in range(world):
            assert int(c[r]) == int(((dest_plain == r) & ~is_hot).sum())
        shuffle_scatter(j, world, [t.data_ptr() for t in sk], [t.data_ptr() for t in sr], [0] * (world + 2))
        j.synchronize()
        hot_rows = sr[world + 1][: int(c[world + 1])].cpu().numpy().view(np.uint32)
        assert np.array_equal(np.sort(hot_rows), np.flatnonzero(is_hot))
        assert np.array_equal(sk[world + 1][: int(c[world + 1])].cpu().numpy(), keys[hot_rows])
        for r in range(world):
            rows = sr[r][: int(c[r])].cpu().numpy().view(np.uint32)
            assert not is_hot[rows].any() and (dest_plain[rows] == r).all()
            assert np.array_equal(sk[r][: int(c[r])].cpu().numpy(), keys[rows])
        shuffle_set_hot(j, None)
        assert np.array_equal(shuffle_count(j, dev, world)[: world + 1], plain[: world + 1])
        shuffle_scatter(j, world, [t.data_ptr() for t in sk[: world + 1]], [t.data_ptr() for t in sr[: world + 1]], [0] * (world + 1))


@pytest.mark.gpu
@pytest.mark.parametrize("dtype,world,null_frac", [(np.int64, 2, 0.0), (np.int64, 8, 0.05), (np.int32, 3, 0.0)])
def test_single_pass_local_scatter_matches_two_pass(dtype, world, null_frac):
    """b200j_shuffle_local_scatter (one read of the keys, regions reserved tile by tile) delivers exactly the rows that the
    two-pass count + scatter delivers to every destination; a region that is too small is an error, not a silent overrun."""
    import torch
    import datafusion_parallelism_b200 as d
    from datafusion_parallelism_b200.exchange import shuffle_count, shuffle_scatter, shuffle_local_scatter
    rng = np.random.default_rng(13)
    kw, n = np.dtype(dtype).itemsize, 300_001
    tdt = torch.int64 if kw == 8 else torch.int32
    keys = rng.integers(-5000, 1 << 30, n).astype(dtype)
    with d.HashJoin("inner", key_widths=(kw,)) as j:
        dev = torch.from_numpy(np.concatenate([keys[:1], keys])).cuda()[1:]          # misaligned start as well
        col = dev
        if null_frac > 0:
            valid = rng.random(n) >= null_frac
            vbits = torch.from_numpy(np.packbits(valid.astype(np.uint8), bitorder="little")).cuda()
            col = d.column(dev, validity=vbits)
        counts = shuffle_count(j, col, world)
        sk = [torch.empty(n, dtype=tdt, device="cuda") for _ in range(world + 1)]
        sr = [torch.empty(n, dtype=torch.int32, device="cuda") for _ in range(world + 1)]
        shuffle_scatter(j, world, [t.data_ptr() for t in sk], [t.data_ptr() for t in sr], [0] * (world + 1))
        j.synchronize()
        cap = int(counts[: world + 1].max()) + 17
        send_k = torch.full(((world + 2) * cap,), -1, dtype=tdt, device="cuda")
        send_r = torch.full(((world + 2) * cap,), -1, dtype=torch.int32, device="cuda")
        got = shuffle_local_scatter(j, col, world, send_k.data_ptr(), send_r.data_ptr(), cap, rowid_base=1000)
        assert np.array_equal(got[: world + 1], counts[: world + 1])
        for dst in range(world + 1):
            m = int(counts[dst])
            want_rows = np.sort(sr[dst][:m].cpu().numpy().view(np.uint32))
            rows = send_r[dst * cap: dst * cap + m].cpu().numpy().view(np.uint32)
            ks = send_k[dst * cap: dst * cap + m].cpu().numpy()
            assert np.array_equal(np.sort(rows) - 1000, want_rows)
            if dst < world:
                assert np.array_equal(ks, keys[(rows - 1000).astype(np.int64)])
            assert (send_r[dst * cap + m: (dst + 1) * cap].cpu().numpy() == -1).all()       # nothing past the reserved rows
        small = max(1, int(counts[:world].min()) // 2)
        with pytest.raises(d.B200JoinError) as e:
            shuffle_local_scatter(j, col, world, send_k.data_ptr(), send_r.data_ptr(), small)
        assert "region too small" in str(e.value)
        assert (send_r[(world + 1) * cap:].cpu().numpy() == -1).all()                       # and no write beyond the regions handed in


@pytest.mark.gpu
def test_local_scatter_narrow_keys():
    """key_out_width = 4: the regions receive 8-byte keys as 32-bit values; a key that does not fit fails the call."""
    import torch
    import datafusion_parallelism_b200 as d
    from datafusion_parallelism_b200.exchange import shuffle_local_scatter
    rng = np.random.default_rng(2)
    n, world = 150_003, 4
    keys = rng.integers(0, 2**32, n).astype(np.int64)
    with d.HashJoin("inner", key_widths=(8,)) as j:
        dev = torch.from_numpy(keys).cuda()
        cap = n
        wide_k = torch.empty((world + 2) * cap, dtype=torch.int64, device="cuda")
        wide_r = torch.empty((world + 2) * cap, dtype=torch.int32, device="cuda")
        c8 = shuffle_local_scatter(j, dev, world, wide_k.data_ptr(), wide_r.data_ptr(), cap)
        nar_k = torch.empty((world + 2) * cap, dtype=torch.int32, device="cuda")
        nar_r = torch.empty((world + 2) * cap, dtype=torch.int32, device="cuda")
        c4 = shuffle_local_scatter(j, dev, world, nar_k.data_ptr(), nar_r.data_ptr(), cap, key_out_width=4)
        assert np.array_equal(c4, c8)
        for dst in range(world):
            m = int(c8[dst])
            rows = nar_r[dst * cap: dst * cap + m].cpu().numpy().view(np.uint32).astype(np.int64)
            got = nar_k[dst * cap: dst * cap + m].cpu().numpy().view(np.uint32).astype(np.int64)
            assert np.array_equal(got, keys[rows])
            assert np.array_equal(np.sort(rows), np.sort(wide_r[dst * cap: dst * cap + m].cpu().numpy().view(np.uint32).astype(np.int64)))
        for bad in (-1, 2**32):
            keys2 = keys.copy()
            keys2[n // 2] = bad
            with pytest.raises(d.B200JoinError) as e:
                shuffle_local_scatter(j, torch.from_numpy(keys2).cuda(), world, nar_k.data_ptr(), nar_r.data_ptr(), cap, key_out_width=4)
            assert "do not fit 32 bits" in str(e.value)
