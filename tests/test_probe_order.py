"""B200J_OUT_PROBE_ORDER: against unique build keys every probe row has at most one match, and the result is ONE array in the
order of the probe batch -- build_idx[i] = the build row probe row i matches, or the NULL build index -- instead of index pairs
(half the bytes back over PCIe; the probe columns need no take). Checked position by position against numpy, for the join
types that honour it, both layouts, host / device / pipelined-host batches and both build-index widths."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

NULL64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def _expected(bk, bvalid, pk, pvalid):
    pos = {int(k): i for i, k in enumerate(bk) if bvalid[i]}
    return np.array([pos.get(int(k), -1) if pvalid[i] else -1 for i, k in enumerate(pk)], dtype=np.int64)


@pytest.mark.parametrize("join_type", ["inner", "left", "right", "full"])
@pytest.mark.parametrize("layout,radix_bits", [("direct", 0), ("radix", 4)])
@pytest.mark.parametrize("mem", ["host", "device", "host_pipelined"])
@pytest.mark.parametrize("u32", [False, True])
def test_probe_order_matches_numpy(join_type, layout, radix_bits, mem, u32):
    import torch
    import pyarrow as pa
    import datafusion_parallelism_b200 as d
    from datafusion_parallelism_b200.join import column
    rng = np.random.default_rng(11)
    nb, npr = 30_000, 200_003
    bk = rng.permutation(nb).astype(np.int64)                       # unique, includes the key value 0
    bmask = rng.random(nb) < 0.03                                    # NULL build keys never match
    pk = rng.integers(-3_000, nb + 3_000, npr).astype(np.int64)     # ~17 % misses
    pmask = rng.random(npr) < 0.04
    want_rows = _expected(bk, ~bmask, pk, ~pmask)
    want = d.OUT_PAIRS | d.OUT_PROBE_ORDER | (d.OUT_BUILD_U32 if u32 else 0) | (0 if mem == "device" else d.OUT_HOST)
    saved = os.environ.get("B200J_HOST_CHUNK_ROWS")
    if mem == "host_pipelined":
        os.environ["B200J_HOST_CHUNK_ROWS"] = "40001"              # 5 ragged chunks through the pipelined path
    try:
        with d.HashJoin(join_type, key_widths=(8,), layout=layout, radix_bits=radix_bits) as j:
            j.build_push(0, [pa.array(bk, mask=bmask)])
            j.build_finish(0)
            if mem == "device":
                valid = torch.from_numpy(np.packbits(~pmask, bitorder="little")).cuda()
                keys = [column(torch.from_numpy(pk).cuda(), validity=valid)]
            else:
                keys = [pa.array(pk, mask=pmask)]
            res = j.probe_batch(0, keys, want=want)
            assert res.probe_order and res.n == npr and res.probe_ptr == 0
            assert res.build_idx_width == (4 if u32 else 8)
            if mem == "device":
                b_, p_ = res.device_arrays()
                assert p_ is None
                got = torch.as_tensor(b_, device="cuda").cpu().numpy().astype(np.int64)
                if u32:
                    got = np.where((got & 0xFFFFFFFF) == 0xFFFFFFFF, -1, got & 0xFFFFFFFF)
            else:
                b, p = res.numpy()
                assert np.array_equal(p, np.arange(npr, dtype=np.uint32))
                got = np.where(b == NULL64, -1, b.astype(np.int64))
            assert res.n_build_null == int((want_rows < 0).sum())
            res.release()
            st = j.stats()
            fin = j.probe_finish(0)
            fb, fp = fin.numpy()
            fin.release()
        assert st["probe_order_results"] == 1
        assert np.array_equal(got, want_rows)
        if join_type in ("left", "full"):                            # the finaliser's unmatched build rows: every build row no probe row hit
            hit = np.zeros(nb, dtype=bool)
            hit[want_rows[want_rows >= 0]] = True
            assert np.array_equal(np.sort(fb.astype(np.int64)), np.flatnonzero(~hit))
        else:
            assert len(fb) == 0
    finally:
        if saved is None:
            os.environ.pop("B200J_HOST_CHUNK_ROWS", None)
        else:
            os.environ["B200J_HOST_CHUNK_ROWS"] = saved


def test_probe_order_is_declined_for_duplicate_build_keys_and_semi_joins():
    import oracle
    import datafusion_parallelism_b200 as d
    from helpers import pairs_multiset
    rng = np.random.default_rng(12)
    bk = rng.integers(0, 5_000, 20_000).astype(np.int64)           # duplicates
    pk = rng.integers(0, 6_000, 50_000).astype(np.int64)
    with d.HashJoin("inner", key_widths=(8,)) as j:
        j.build_push(0, [bk])
        j.build_finish(0)
        res = j.probe_batch(0, [pk], want=d.OUT_PAIRS | d.OUT_HOST | d.OUT_PROBE_ORDER)
        assert not res.probe_order and res.probe_ptr != 0
        b, p = res.numpy()
        res.release()
    ob, op = oracle.OracleJoin("inner", [bk]).probe_batch([pk], len(pk))
    g, w = pairs_multiset(b, p), pairs_multiset(ob, op)
    assert np.array_equal(g[0], w[0]) and np.array_equal(g[1], w[1])
    ub = rng.permutation(5_000).astype(np.int64)
    with d.HashJoin("right_semi", key_widths=(8,)) as j:
        j.build_push(0, [ub])
        j.build_finish(0)
        res = j.probe_batch(0, [pk], want=d.OUT_PAIRS | d.OUT_HOST | d.OUT_PROBE_ORDER)
        assert not res.probe_order
        _, p = res.numpy()
        res.release()
    assert np.array_equal(np.sort(p), np.flatnonzero(pk < 5_000).astype(np.uint32))


def test_gather_of_a_probe_order_result():
    """b200j_gather takes a probe-order result with probe_batch = NULL: the build columns come out in the order of the probe
    batch (NULL rows where a probe row has no match); handing it a probe batch is refused."""
    import ctypes as C
    import datafusion_parallelism_b200 as d
    rng = np.random.default_rng(13)
    nb, npr = 8_000, 30_000
    bk = rng.permutation(nb).astype(np.int64)
    bpay = (bk * 5 + 2).astype(np.int64)
    pk = rng.integers(0, nb + 2_000, npr).astype(np.int64)
    with d.HashJoin("right", key_widths=(8,), build_payload_types=(8,), layout="radix", radix_bits=3) as j:
        j.build_push(0, [bk], [bpay])
        j.build_finish(0)
        res = j.probe_batch(0, [pk], want=d.OUT_PAIRS | d.OUT_PROBE_ORDER)          # device result
        assert res.probe_order
        with pytest.raises(d.B200JoinError):
            j.gather(res, [pk])
        g = j.gather(res, None, want=d.OUT_HOST)
        n = g.n_rows
        cols = [g.column_buffers(i) for i in range(2)]
        key = np.frombuffer(C.string_at(cols[0]["values"], n * 8), dtype=np.int64)
        pay = np.frombuffer(C.string_at(cols[1]["values"], n * 8), dtype=np.int64)
        valid = np.unpackbits(np.frombuffer(C.string_at(cols[0]["validity"], (n + 31) // 32 * 4), dtype=np.uint8), bitorder="little")[:n].astype(bool)
        g.release()
        res.release()
    assert n == npr and np.array_equal(valid, pk < nb)
    assert np.array_equal(key[valid], pk[valid]) and np.array_equal(pay[valid], pk[valid] * 5 + 2)


@pytest.mark.parametrize("mem", ["host", "device"])
def test_probe_order_of_an_empty_and_a_one_row_batch(mem):
    import torch
    import datafusion_parallelism_b200 as d
    bk = np.arange(1, 1001, dtype=np.int64)
    with d.HashJoin("inner", key_widths=(8,)) as j:
        j.build_push(0, [bk])
        j.build_finish(0)
        for pk in (np.zeros(0, dtype=np.int64), np.array([7], dtype=np.int64), np.array([5000], dtype=np.int64)):
            keys = [torch.from_numpy(pk).cuda()] if mem == "device" else [pk]
            res = j.probe_batch(0, keys, want=d.OUT_PAIRS | d.OUT_PROBE_ORDER | (0 if mem == "device" else d.OUT_HOST))
            assert res.probe_order and res.n == len(pk) and res.probe_ptr == 0
            if mem == "host" and len(pk):
                b, p = res.numpy()
                assert (int(b[0]) == 6) if pk[0] == 7 else (b[0] == NULL64)
            elif len(pk):
                b_, _ = res.device_arrays()
                got = int(torch.as_tensor(b_, device="cuda").cpu().numpy()[0])
                assert got == (6 if pk[0] == 7 else -1)
            res.release()
