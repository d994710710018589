"""GPU parity of the output gather (a6, take_multiple_record_batch, src/shared/shared.rs:83-92): the gathered output
columns, compared as an order-normalised multiset of full rows with the oracle's whole-join output — the comparison
BASELINE.json's north_star names."""
import numpy as np
import pyarrow as pa
import pytest

import oracle
from helpers import rows_multiset

pytestmark = pytest.mark.gpu

JOIN_TYPES = ["inner", "left", "right", "full", "left_semi", "right_semi", "left_anti", "right_anti"]


def _tables(rng, n_build, n_probe, key_range, null_frac):
    def mk(n, prefix):
        k = rng.integers(0, key_range, n).astype(np.int32)
        kmask = rng.random(n) < null_frac
        v64 = rng.integers(-10**12, 10**12, n).astype(np.int64)
        vmask = rng.random(n) < 0.1
        strs = [None if rng.random() < 0.1 else prefix + "x" * int(rng.integers(0, 12)) + str(int(rng.integers(0, 1000))) for _ in range(n)]
        return pa.table({"id": pa.array(k, type=pa.int32(), mask=kmask if null_frac > 0 else None),
                         "v": pa.array(v64, type=pa.int64(), mask=vmask),
                         "s": pa.array(strs, type=pa.string())})
    return mk(n_build, "b"), mk(n_probe, "p")


@pytest.mark.parametrize("join_type", JOIN_TYPES)
@pytest.mark.parametrize("layout,radix_bits", [("direct", 0), ("radix", 4)])
@pytest.mark.parametrize("null_frac", [0.0, 0.08])
def test_gathered_rows_match_oracle(join_type, layout, radix_bits, null_frac):
    import datafusion_parallelism_b200 as d
    rng = np.random.default_rng(17)
    build, probe = _tables(rng, 3000, 7000, 2500, null_frac)
    probe_batches = [probe.slice(0, 4000), probe.slice(4000)]
    want = oracle.join_tables(build, probe_batches, [("id", "id")], join_type)
    left_only, right_only = join_type in ("left_semi", "left_anti"), join_type in ("right_semi", "right_anti")
    types = ([] if right_only else [pa.int32(), pa.int64(), pa.string()]) + ([] if left_only else [pa.int32(), pa.int64(), pa.string()])
    pieces = []
    with d.HashJoin(join_type, key_widths=(4,), build_payload_types=(8, d.T_UTF8), probe_payload_types=(8, d.T_UTF8),
                    layout=layout, radix_bits=radix_bits) as j:
        j.build_push(0, [build.column("id")], [build.column("v"), build.column("s")])
        j.build_finish(0)
        for pb in probe_batches:
            pb = pb.combine_chunks()
            res = j.probe_batch(0, [pb.column("id")], want=d.OUT_PAIRS)          # pairs stay on the device
            g = j.gather(res, [pb.column("id")], [pb.column("v"), pb.column("s")])
            if not left_only:
                pieces.append(pa.Table.from_arrays(g.to_arrow(types), names=want.column_names))
            g.release()
            res.release()
        fin = j.probe_finish(0, want=d.OUT_PAIRS | d.OUT_HOST)                   # host pairs: staged by the gather
        if fin.n or not pieces:
            g = j.gather(fin, None)
            pieces.append(pa.Table.from_arrays(g.to_arrow(types), names=want.column_names))
            g.release()
        fin.release()
    got = pa.concat_tables(pieces)
    assert got.num_rows == want.num_rows
    assert rows_multiset(got) == rows_multiset(want)


def test_gather_device_probe_batch_and_wide_column():
    """Device-resident probe batch, 16-byte payload (Decimal128-sized), gathered to device then read back."""
    import torch
    import datafusion_parallelism_b200 as d
    rng = np.random.default_rng(3)
    nb, npr = 5000, 20000
    bk = rng.permutation(nb).astype(np.int64)
    bw = rng.integers(0, 2**62, (nb, 2)).astype(np.int64)                          # 16 B per row
    pk = rng.integers(0, nb + 500, npr).astype(np.int64)
    with d.HashJoin("inner", key_widths=(8,), build_payload_types=(16,), layout="radix", radix_bits=3) as j:
        bw16 = np.ascontiguousarray(bw).view([("a", "<i8"), ("b", "<i8")]).reshape(nb)      # one 16-byte value per row
        j.build_push(0, [bk], [bw16])
        j.build_finish(0)
        dpk = torch.from_numpy(pk).cuda()
        res = j.probe_batch(0, [dpk], want=d.OUT_PAIRS)
        g = j.gather(res, [dpk], want=d.OUT_HOST)
        cols = [g.column_buffers(i) for i in range(g.n_cols)]
        n = g.n_rows
        import ctypes as C
        gbk = np.frombuffer(C.string_at(cols[0]["values"], n * 8), dtype=np.int64)
        gbw = np.frombuffer(C.string_at(cols[1]["values"], n * 16), dtype=np.int64).reshape(n, 2)
        gpk = np.frombuffer(C.string_at(cols[2]["values"], n * 8), dtype=np.int64)
        g.release()
        res.release()
    assert n == int((pk < nb).sum())
    assert np.array_equal(gbk, gpk)                                                # key equality on every output row
    pos = np.empty(nb, dtype=np.int64)
    pos[bk] = np.arange(nb)
    assert np.array_equal(gbw, bw[pos[gpk]])
    assert sorted(gpk.tolist()) == sorted(pk[pk < nb].tolist())


def _gather_columns(env, join_type, bk, bcols, btypes, pk_dev, pk_valid, ppay_dev, want_pairs, pieces=None):
    """Build + probe + gather on a fresh handle created under `env`; returns (stats, [(values bytes, validity bits | None, null_count)])."""
    import ctypes as C
    import os
    import datafusion_parallelism_b200 as d
    from datafusion_parallelism_b200.join import column
    saved = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
        j = d.HashJoin(join_type, key_widths=(8,), build_payload_types=btypes, probe_payload_types=(8,), layout="radix", radix_bits=4)
    finally:
        for k, v in saved.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    with j:
        j.build_push(0, [bk], bcols)
        j.build_finish(0)
        if pieces is None:
            keys = [pk_dev] if pk_valid is None else [column(pk_dev, validity=pk_valid)]
            res = j.probe_batch(0, keys, want=want_pairs)
        else:
            keys = [pk_dev]
            res = j.probe_pieces(0, pieces, want=want_pairs)
        g = j.gather(res, keys, [ppay_dev], want=d.OUT_HOST)
        n, out = g.n_rows, []
        widths = [8] + list(btypes) + [8, 8]
        assert g.n_cols == len(widths)
        for i, w in enumerate(widths):
            c = g.column_buffers(i)
            vals = np.frombuffer(C.string_at(c["values"], n * w), dtype=np.uint8).reshape(n, w).copy()
            bits = None
            if c["validity"]:
                bits = np.unpackbits(np.frombuffer(C.string_at(c["validity"], ((n + 31) // 32) * 4), dtype=np.uint8), bitorder="little")
                assert not bits[n:].any(), "bits past the last row must be zero"
                bits = bits[:n].astype(bool)
                vals[~bits] = 0
            out.append((vals, bits, c["null_count"]))
        st = j.stats()
        g.release()
        res.release()
    return st, out


@pytest.mark.parametrize("join_type,miss", [("inner", False), ("right", True), ("full", True)])
@pytest.mark.parametrize("u32", [False, True])
@pytest.mark.parametrize("shape", ["batch", "misaligned", "pieces"])
@pytest.mark.parametrize("hint", [False, True])
def test_partition_order_gather_equals_plain_gather(join_type, miss, u32, shape, hint):
    """The build side of a dense radix result is gathered partition by partition (gather_seg_multi_kernel) and both key columns
    come out of one pass (gather_keys_both_kernel): same columns, bit for bit, as the plain row-order kernels, and as numpy."""
    import torch
    import datafusion_parallelism_b200 as d
    rng = np.random.default_rng(23)
    nb, npr = 60_000, 150_003                                     # 5 blocks of 32768 rows, the last one ragged
    bk = rng.permutation(nb).astype(np.int64) + 1
    b64 = rng.integers(-2**62, 2**62, nb).astype(np.int64)
    b64_mask = rng.random(nb) < 0.07
    b32 = rng.integers(-2**31, 2**31 - 1, nb).astype(np.int32)
    b16 = rng.integers(0, 2**62, (nb, 2)).astype(np.int64).view([("a", "<i8"), ("b", "<i8")]).reshape(nb)
    b8 = rng.integers(0, 255, nb).astype(np.uint8)
    b2 = rng.integers(0, 65535, nb).astype(np.uint16)
    bcols = [pa.array(b64, mask=b64_mask), b32, b16, b8, b2]
    btypes = (8, 4, 16, 1, 2)
    pk = rng.integers(1, nb + (nb // 5 if miss else 1), npr).astype(np.int64)
    ppay = rng.integers(0, 2**40, npr).astype(np.int64)
    store = torch.empty(npr + 1, dtype=torch.int64, device="cuda")
    pk_dev = store[1:] if shape == "misaligned" else store[:npr]     # 8 bytes off a 16-byte boundary
    pk_dev.copy_(torch.from_numpy(pk))
    ppay_dev = torch.from_numpy(ppay).cuda()
    pk_valid = None
    if miss and shape != "pieces":                                # NULL probe keys: the rows of the stream's last segment per block
        vmask = rng.random(npr) < 0.05
        pk_valid = torch.from_numpy(np.packbits(~vmask, bitorder="little")).cuda()
    else:
        vmask = np.zeros(npr, dtype=bool)
    pieces = None
    if shape == "pieces":
        cut = 70_001
        pieces = [(pk_dev[:cut].data_ptr(), cut), (pk_dev[cut:].data_ptr(), npr - cut)]
    want = d.OUT_PAIRS | (d.OUT_BUILD_U32 if u32 else 0)
    # hint: B200J_OUT_GATHER_HINT -- key columns from the partitioned key stream (gather_keys_stream_kernel, batches without NULL
    # keys), fixed-width payload from the slot-ordered copies (slot_order_kernel) at the slot the probe recorded for every pair
    st_new, new = _gather_columns({}, join_type, bk, bcols, btypes, pk_dev, pk_valid, ppay_dev, want | (d.OUT_GATHER_HINT if hint else 0), pieces)
    st_old, old = _gather_columns({"B200J_GATHER_SEG": "0", "B200J_KEY_VIA_PROBE": "0"}, join_type, bk, bcols, btypes, pk_dev, pk_valid,
                                  ppay_dev, want, pieces)
    assert st_new["gather_seg_launches"] >= 1 and st_new["seg_probe_launches"] >= 1, st_new
    assert st_new["gather_hint_launches"] == (1 if hint else 0), st_new
    assert st_old["gather_seg_launches"] == 0 and st_old["gather_hint_launches"] == 0, st_old
    assert len(new) == len(old)
    for c, ((v1, m1, z1), (v2, m2, z2)) in enumerate(zip(new, old)):
        assert z1 == z2, (c, z1, z2)
        assert (m1 is None) == (m2 is None), c

    def sorted_rows(cols):
        # the order of the rows inside a segment of the partitioned stream differs from run to run (shared-memory atomics rank
        # them): whole rows -- every column's bytes and validity -- are compared as sorted multisets
        n_ = cols[0][0].shape[0]
        parts = []
        for v, m, _ in cols:
            parts.append(v)
            parts.append((m if m is not None else np.ones(n_, dtype=bool)).astype(np.uint8).reshape(n_, 1))
        mat = np.ascontiguousarray(np.concatenate(parts, axis=1))
        return np.sort(mat.view([("r", np.void, mat.shape[1])]).reshape(n_), axis=0)

    assert np.array_equal(sorted_rows(new), sorted_rows(old))
    # ... and against numpy on the rows themselves
    n = new[0][0].shape[0]
    gpk = new[-2][0].view(np.int64).reshape(n)
    gbk = new[0][0].view(np.int64).reshape(n)
    hit = new[0][1] if new[0][1] is not None else np.ones(n, dtype=bool)
    pos = np.zeros(nb + 1 + nb // 5 + 1, dtype=np.int64)
    pos[bk] = np.arange(nb)
    assert np.array_equal(gbk[hit], gpk[hit])
    rows = pos[gpk[hit]]
    assert np.array_equal(new[2][0].view(np.int32).reshape(n)[hit], b32[rows])
    assert np.array_equal(new[3][0][hit], np.ascontiguousarray(b16).view(np.uint8).reshape(nb, 16)[rows])
    assert np.array_equal(new[4][0].reshape(n)[hit], b8[rows])
    assert np.array_equal(new[5][0].view(np.uint16).reshape(n)[hit], b2[rows])
    v64 = new[1][0].view(np.int64).reshape(n)
    m64 = new[1][1] if new[1][1] is not None else np.ones(n, dtype=bool)
    assert np.array_equal(m64[hit], ~b64_mask[rows])
    assert np.array_equal(v64[hit & m64], b64[rows][m64[hit]])
    if join_type == "inner":
        assert n == int(((pk <= nb) & ~vmask).sum())
    elif join_type == "right":
        assert n == npr


def test_hinted_gather_of_a_join_without_payload():
    """B200J_OUT_GATHER_HINT on a join that has only its key column: the result keeps the partitioned key stream (no slot array:
    there is no payload to read through it) and both key columns of the output are copied from it."""
    import ctypes as C
    import torch
    import datafusion_parallelism_b200 as d
    rng = np.random.default_rng(5)
    nb, npr = 40_000, 100_001
    bk = rng.permutation(nb).astype(np.int64) * 3 + 1
    pk = bk[rng.integers(0, nb, npr)]
    pk_dev = torch.from_numpy(pk).cuda()
    with d.HashJoin("inner", key_widths=(8,), layout="radix", radix_bits=3) as j:
        j.build_push(0, [bk])
        j.build_finish(0)
        res = j.probe_batch(0, [pk_dev], want=d.OUT_PAIRS | d.OUT_GATHER_HINT)
        b_, p_ = res.device_arrays()
        pidx = torch.as_tensor(p_, device="cuda").cpu().numpy().astype(np.int64) & 0xFFFFFFFF
        bidx = torch.as_tensor(b_, device="cuda").cpu().numpy().astype(np.int64)
        g = j.gather(res, [pk_dev], want=d.OUT_HOST)
        n = g.n_rows
        cols = [g.column_buffers(i) for i in range(g.n_cols)]
        gb = np.frombuffer(C.string_at(cols[0]["values"], n * 8), dtype=np.int64).copy()
        gp = np.frombuffer(C.string_at(cols[1]["values"], n * 8), dtype=np.int64).copy()
        st = j.stats()
        g.release()
        res.release()
    assert n == npr and g.n_cols == 2
    assert all(c["null_count"] == 0 and not c["validity"] for c in cols)
    assert st["gather_hint_launches"] == 1 and st["seg_probe_launches"] == 1, st
    assert np.array_equal(gp, pk[pidx]) and np.array_equal(gb, bk[bidx]) and np.array_equal(gb, gp)
