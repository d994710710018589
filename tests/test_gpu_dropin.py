"""GPU tests of the drop-in claim itself: the reference's own golden scenarios (src/lib.rs) through the CUDA path
(HashJoin + gather + the JoinFilter round trip, compared as ROWS with the literal expected batches), the call pattern of
N DataFusion partitions on N host threads (parallel_hash_join.rs:144-167), and the borrowed-buffer / bad-index rules of
include/b200join.h."""
import threading

import numpy as np
import pyarrow as pa
import pyarrow.compute as pc
import pytest

import oracle
from helpers import load_golden, rows_multiset, table_from_dict

pytestmark = pytest.mark.gpu

G = load_golden()


@pytest.fixture(scope="module")
def pkg():
    import datafusion_parallelism_b200 as d
    assert d.device_count() >= 1, "no CUDA device: -m gpu tests must run on the B200 box"
    return d


def _width_or_utf8(pkg, typ):
    return pkg.T_UTF8 if (pa.types.is_string(typ) or pa.types.is_binary(typ)) else typ.bit_width // 8


def gpu_join_tables(pkg, build: pa.Table, probe_batches, on, join_type, filter_expr=None, layout="auto", radix_bits=0,
                    build_batch_rows=None):
    """The whole join through the C ABI the way the Rust operator would drive it: build batches pushed, every probe batch
    probed (pairs stay on the device unless a JoinFilter has to see them) and gathered, then the finaliser batch.
    Returns a table with the oracle's column naming (l.<build columns in schema order>, r.<probe columns>)."""
    bkey, pkey = [b for b, _ in on], [p for _, p in on]
    bpay = [n for n in build.column_names if n not in bkey]
    probe_schema = probe_batches[0].schema
    ppay = [n for n in probe_schema.names if n not in pkey]
    left_only, right_only = join_type in ("left_semi", "left_anti"), join_type in ("right_semi", "right_anti")
    out_names, out_types = [], []
    if not right_only:
        out_names += [f"l.{n}" for n in bkey + bpay]
        out_types += [build.schema.field(n).type for n in bkey + bpay]
    if not left_only:
        out_names += [f"r.{n}" for n in pkey + ppay]
        out_types += [probe_schema.field(n).type for n in pkey + ppay]
    want_order = ([f"l.{n}" for n in build.column_names] if not right_only else []) + \
                 ([f"r.{n}" for n in probe_schema.names] if not left_only else [])
    pieces = []
    with pkg.HashJoin(join_type, key_widths=tuple(_width_or_utf8(pkg, build.schema.field(n).type) for n in bkey),
                      build_payload_types=tuple(_width_or_utf8(pkg, build.schema.field(n).type) for n in bpay),
                      probe_payload_types=tuple(_width_or_utf8(pkg, probe_schema.field(n).type) for n in ppay),
                      layout=layout, radix_bits=radix_bits) as j:
        step = build_batch_rows or max(build.num_rows, 1)
        for lo in range(0, max(build.num_rows, 1), step):
            bb = build.slice(lo, step).combine_chunks()
            j.build_push(0, [bb.column(n) for n in bkey], [bb.column(n) for n in bpay])
        j.build_finish(0)
        for pb in probe_batches:
            pb = pb.combine_chunks()
            keys, pay = [pb.column(n) for n in pkey], [pb.column(n) for n in ppay]
            if filter_expr is None:
                res = j.probe_batch(0, keys, want=pkg.OUT_PAIRS)
            else:
                # JoinFilter stays with the caller (apply_join_filter_to_indices, datafusion_private.rs:295-328)
                assert filter_expr == "l.value != r.value"
                m = j.probe_batch(0, keys, want=pkg.OUT_PAIRS | pkg.OUT_HOST | pkg.OUT_MATCHES_ONLY)
                b, p = m.numpy()
                m.release()
                lv = build.column("value").take(pa.array(b.astype(np.int64)))
                rv = pb.column("value").take(pa.array(p.astype(np.int64)))
                keep = np.asarray(pc.fill_null(pc.not_equal(lv, rv), False)) if len(b) else np.zeros(0, dtype=bool)
                res = j.apply_pairs(0, pb.num_rows, b[keep], p[keep], want=pkg.OUT_PAIRS | pkg.OUT_HOST)
            if not left_only:
                g = j.gather(res, keys, pay)
                pieces.append(pa.Table.from_arrays(g.to_arrow(out_types), names=out_names))
                g.release()
            res.release()
        fin = j.probe_finish(0, want=pkg.OUT_PAIRS | pkg.OUT_HOST)
        if fin.n or not pieces:
            g = j.gather(fin, None)
            pieces.append(pa.Table.from_arrays(g.to_arrow(out_types), names=out_names))
            g.release()
        fin.release()
    return pa.concat_tables(pieces).select(want_order)


@pytest.mark.parametrize("sc", G["scenarios"], ids=[s["name"] for s in G["scenarios"]])
@pytest.mark.parametrize("probe_split", [None, 1, 2])
@pytest.mark.parametrize("layout,radix_bits", [("direct", 0), ("radix", 2)])
def test_lib_rs_scenarios_through_cuda(pkg, sc, probe_split, layout, radix_bits):
    """The ten src/lib.rs scenarios (Int32 key with NULLs, Utf8 payload, all join types, the post-join JoinFilter): output
    ROWS of the CUDA path against the reference's literal expected batches."""
    left, right = table_from_dict(sc["left"]), table_from_dict(sc["right"])
    batches = [right] if probe_split is None else [right.slice(i, probe_split) for i in range(0, right.num_rows, probe_split)]
    out = gpu_join_tables(pkg, left, batches, sc["on"], sc["join_type"], sc["filter"], layout, radix_bits)
    if sc.get("post_filter"):
        assert sc["post_filter"] == "r.id is not null"
        out = out.filter(pc.is_valid(out.column("r.id")))
    assert rows_multiset(out) == rows_multiset(sc["expected"])
    # and the same rows as the oracle's whole-join output, column by column
    want = oracle.join_tables(left, batches, sc["on"], sc["join_type"], None if sc["filter"] is None else
                              (lambda b, p, bi, pi: np.asarray(pc.fill_null(pc.not_equal(
                                  b.column("value").take(pa.array(bi.astype(np.int64))),
                                  p.column("value").take(pa.array(pi.astype(np.int64)))), False))))
    if sc.get("post_filter"):
        want = want.filter(pc.is_valid(want.column("r.id")))
    assert rows_multiset(out) == rows_multiset(want)


def test_inner_join_no_filter_chain_through_cuda(pkg):
    """inner_join_no_filter (src/lib.rs:67-132): a 1024-row table joined four times with a small table, 64-row batches on
    both sides; every join's output is the next join's build side."""
    sc = G["chain_scenario"]
    n = sc["batch_size"] * sc["batches"]
    ids = pa.array(np.arange(n, dtype=np.int32))
    acc = pa.Table.from_arrays([ids, ids, ids, ids, pa.array(["hello"] * n)], names=["id1", "id2", "id3", "id4", "value"])
    small = pa.Table.from_arrays([ids, pa.array(["world"] * n)], names=["id", "value"])
    for k, (lk, rk) in enumerate(sc["joins"]):
        probe_batches = [small.slice(i, sc["batch_size"]) for i in range(0, n, sc["batch_size"])]
        acc = gpu_join_tables(pkg, acc, probe_batches, [(lk, rk)], "inner", build_batch_rows=sc["batch_size"])
        acc = acc.rename_columns([c[2:] if c.startswith("l.") else f"s{k}.{c[2:]}" for c in acc.column_names])
    assert acc.num_rows == 1024 and acc.num_columns == 13
    order = np.argsort(np.asarray(acc.column(0)))
    for ci in (0, 1, 2, 3, 5, 7, 9, 11):
        assert np.array_equal(np.asarray(acc.column(ci))[order], np.arange(1024))
    assert set(acc.column(4).to_pylist()) == {"hello"}
    for ci in (6, 8, 10, 12):
        assert set(acc.column(ci).to_pylist()) == {"world"}


@pytest.mark.parametrize("join_type", ["inner", "left", "full", "right_anti", "left_semi"])
@pytest.mark.parametrize("layout,radix_bits", [("direct", 0), ("radix", 4)])
def test_partitions_on_concurrent_host_threads(pkg, join_type, layout, radix_bits):
    """What ParallelHashJoin::execute does from N tokio workers (parallel_hash_join.rs:144-167): every partition pushes its
    build batches, finishes its build stream, probes its own probe batches and finishes its probe stream -- all N
    partitions at once, on N host threads, against ONE shared handle. Results are compared per probe batch with the
    oracle; exactly one partition (the last to finish) gets the finaliser rows."""
    n_part, per_part_build, per_part_probe = 6, 5, 7
    rng = np.random.default_rng(hash((join_type, layout, "mt")) & 0xFFFFFFFF)
    build_batches = [[rng.integers(0, 4000, int(rng.integers(1, 900))).astype(np.int64) for _ in range(per_part_build)] for _ in range(n_part)]
    probe_batches = [[rng.integers(-200, 4400, int(rng.integers(1, 3000))).astype(np.int64) for _ in range(per_part_probe)] for _ in range(n_part)]
    results = [[None] * per_part_probe for _ in range(n_part)]
    fins = [None] * n_part
    errors = []
    build_barrier = threading.Barrier(n_part)
    with pkg.HashJoin(join_type, key_widths=(8,), n_partitions=n_part, layout=layout, radix_bits=radix_bits) as j:
        def worker(part):
            try:
                for bb in build_batches[part]:
                    j.build_push(part, [bb])
                j.build_finish(part)
                build_barrier.wait()                  # the reference's probe side waits for the shared build as well
                for i, pb in enumerate(probe_batches[part]):
                    r = j.probe_batch(part, [pb])
                    results[part][i] = r.numpy()
                    r.release()
                r = j.probe_finish(part)
                fins[part] = r.numpy()
                r.release()
            except BaseException as e:           # noqa: BLE001 - reported below
                errors.append((part, repr(e)))
                try:
                    build_barrier.abort()
                except Exception:
                    pass
        threads = [threading.Thread(target=worker, args=(p,)) for p in range(n_part)]
        for t in threads:
            t.start()
        for t in threads:
            t.join(timeout=120)
        assert not errors, errors
        assert all(not t.is_alive() for t in threads)
        stats = j.stats()
    # Row numbering = arrival order, which is a race between the threads: compare KEYS, not row ids. For every probe batch
    # the multiset of (build key, probe row) must equal the oracle's, whatever order the build batches arrived in.
    all_build = np.concatenate([b for part in build_batches for b in part])
    assert stats["build_rows"] == len(all_build)
    oj = oracle.OracleJoin(join_type, [all_build])
    NULLB = np.uint64(0xFFFFFFFFFFFFFFFF)

    n_fin = 0
    for part in range(n_part):
        for i, pb in enumerate(probe_batches[part]):
            wb, wp = oj.probe_batch([pb], len(pb))
            gb, gp = results[part][i]
            assert len(gb) == len(wb), (part, i)
            matched = gb != NULLB
            if join_type not in ("right_anti", "left_semi"):
                # counts per probe row are arrival-order independent
                assert np.array_equal(np.sort(gp[matched]), np.sort(wp[wb != NULLB]))
                assert np.array_equal(np.sort(gp[~matched]), np.sort(wp[wb == NULLB]))
            else:
                assert np.array_equal(np.sort(gp), np.sort(wp))
        n_fin += 1 if len(fins[part][0]) or len(fins[part][1]) else 0
    wfb, wfp = oj.finish()
    total_fin = sum(len(f[0]) for f in fins)
    assert total_fin == len(wfb)
    assert n_fin <= 1                                  # only the last partition emits the finaliser


def test_pair_build_keys_with_concurrent_partitions(pkg):
    """Same concurrent call pattern, but with the pairs checked through the gather: (build key, probe key) of every output
    row must be equal -- this holds for any arrival order of the build batches."""
    n_part = 4
    rng = np.random.default_rng(99)
    build_batches = [[rng.permutation(np.arange(p * 10_000 + q * 2_000, p * 10_000 + (q + 1) * 2_000)).astype(np.int64) for q in range(5)] for p in range(n_part)]
    probe_batches = [[rng.integers(0, n_part * 10_000, 20_000).astype(np.int64) for _ in range(3)] for _ in range(n_part)]
    errors, out = [], []
    lock = threading.Lock()
    barrier = threading.Barrier(n_part)
    with pkg.HashJoin("inner", key_widths=(8,), n_partitions=n_part) as j:
        def worker(part):
            try:
                for bb in build_batches[part]:
                    j.build_push(part, [bb])
                j.build_finish(part)
                barrier.wait()
                for pb in probe_batches[part]:
                    r = j.probe_batch(part, [pb], want=pkg.OUT_PAIRS)
                    g = j.gather(r, [pb])
                    cols = g.to_arrow([pa.int64(), pa.int64()])
                    with lock:
                        out.append((len(pb), cols[0].to_numpy(), cols[1].to_numpy()))
                    g.release()
                    r.release()
                j.probe_finish(part).release()
            except BaseException as e:           # noqa: BLE001
                errors.append(repr(e))
                try:
                    barrier.abort()
                except Exception:
                    pass
        ts = [threading.Thread(target=worker, args=(p,)) for p in range(n_part)]
        [t.start() for t in ts]
        [t.join(timeout=120) for t in ts]
    assert not errors, errors
    assert len(out) == n_part * 3
    for n_rows, bkeys, pkeys in out:
        assert len(bkeys) == n_rows                    # unique build keys covering the probe range: 1:1
        assert np.array_equal(bkeys, pkeys)


def test_device_build_batch_is_borrowed_for_the_call_only(pkg):
    """include/b200join.h: input buffers are borrowed for the duration of the call. A device batch overwritten right after
    build_push returns (timing off, so nothing else synchronises) must not change the build side."""
    import torch
    n = 4_000_000
    keys = np.arange(1, n + 1, dtype=np.int64)
    side = torch.cuda.Stream()
    with pkg.HashJoin("inner", key_widths=(8,), layout="direct") as j:
        for lo in range(0, n, n // 4):
            t = torch.from_numpy(keys[lo:lo + n // 4]).cuda()
            torch.cuda.synchronize()
            j.build_push(0, [t])
            with torch.cuda.stream(side):            # a different stream than the handle's: no implicit ordering
                t.fill_(-7)
        j.build_finish(0)
        torch.cuda.synchronize()
        probe = torch.from_numpy(np.array([1, n // 2, n, -7, n + 1], dtype=np.int64)).cuda()
        r = j.probe_batch(0, [probe], want=pkg.OUT_PAIRS | pkg.OUT_HOST)
        b, p = r.numpy()
        r.release()
    o = np.argsort(p)
    assert p[o].tolist() == [0, 1, 2] and b[o].tolist() == [0, n // 2 - 1, n - 1]


def test_apply_pairs_rejects_out_of_range_indices(pkg):
    bk = np.arange(100, dtype=np.int64)
    for jt in ("left", "right", "full"):
        with pkg.HashJoin(jt, key_widths=(8,)) as j:
            j.build_push(0, [bk])
            j.build_finish(0)
            for b, p in (([100], [0]), ([0], [50]), ([2**40], [1])):
                with pytest.raises(pkg.B200JoinError) as e:
                    j.apply_pairs(0, 50, np.array(b, dtype=np.uint64), np.array(p, dtype=np.uint32))
                assert e.value.code == -1
            # NULL build marker and in-range pairs are fine afterwards
            r = j.apply_pairs(0, 50, np.array([3, 99], dtype=np.uint64), np.array([0, 49], dtype=np.uint32))
            assert r.n >= 2
            r.release()
            j.trim()


@pytest.mark.parametrize("join_type", ["inner", "left", "right", "full", "left_anti", "right_semi"])
@pytest.mark.parametrize("layout,radix_bits", [("direct", 0), ("radix", 4)])
def test_build_indices_as_u32(pkg, monkeypatch, join_type, layout, radix_bits):
    """B200J_OUT_BUILD_U32: same pairs with 32-bit build indices (NULL side = UINT32_MAX), from the device-result path, the
    host-result path, the pipelined host path (many ragged chunks) and the finaliser; the gather accepts them as they are."""
    import torch
    monkeypatch.setenv("B200J_HOST_CHUNK_ROWS", "4099")
    rng = np.random.default_rng(hash((join_type, layout, "u32")) & 0xFFFFFFFF)
    for unique in (True, False):
        nb = 30_000
        bk = (rng.permutation(nb * 2)[:nb] if unique else rng.integers(0, nb // 3, nb)).astype(np.int64)
        pk = rng.integers(-100, nb * 2 + 100, 50_001).astype(np.int64)
        bpay = rng.integers(0, 1 << 40, nb).astype(np.int64)
        oj = oracle.OracleJoin(join_type, [bk])
        want = oj.probe_batch([pk], len(pk))
        want_fin = oj.finish()
        wb, wp = want
        with pkg.HashJoin(join_type, key_widths=(8,), build_payload_types=(8,), layout=layout, radix_bits=radix_bits) as j:
            j.build_push(0, [bk], [bpay])
            j.build_finish(0)
            flags = pkg.OUT_PAIRS | pkg.OUT_BUILD_U32
            # host keys -> pinned host pairs through the pipelined path
            r = j.probe_batch(0, [pk], want=flags | pkg.OUT_HOST | pkg.OUT_MATCHES_ONLY)
            assert r.build_idx_width == 4 or r.n == 0
            hb, hp_ = r.numpy()                                           # widened by the veneer
            r.release()
            mb, mp = oracle.OracleJoin("inner", [bk]).probe_batch([pk], len(pk))      # matches only = the inner join's pairs
            o1, o2 = np.lexsort((hb, hp_)), np.lexsort((mb, mp))
            assert np.array_equal(hb[o1], mb[o2]) and np.array_equal(hp_[o1], mp[o2])
            dpk = torch.from_numpy(pk).cuda()
            r = j.probe_batch(0, [dpk], want=flags)                      # device keys -> device pairs
            assert r.mem == pkg.MEM_DEVICE
            b_, p_ = r.device_arrays()
            if r.n and b_ is not None:
                assert r.build_idx_width == 4
                gb = torch.as_tensor(b_, device="cuda").cpu().numpy().view(np.uint32)
                gp = torch.as_tensor(p_, device="cuda").cpu().numpy().view(np.uint32)
                gb64 = np.where(gb == np.uint32(0xFFFFFFFF), np.uint64(0xFFFFFFFFFFFFFFFF), gb.astype(np.uint64))
                o1, o2 = np.lexsort((gb64, gp)), np.lexsort((wb, wp))
                assert np.array_equal(gb64[o1], wb[o2]) and np.array_equal(gp[o1], wp[o2])
                if join_type in ("inner", "left", "right", "full"):
                    g = j.gather(r, [dpk], want=pkg.OUT_HOST)            # u32 build indices straight into the gather
                    cols = g.to_arrow([pa.int64(), pa.int64(), pa.int64()])
                    gk, gpay, gpk = (c.to_numpy(zero_copy_only=False) for c in cols)
                    g.release()
                    m = gb != np.uint32(0xFFFFFFFF)
                    assert np.array_equal(gk[m].astype(np.int64), bk[gb[m].astype(np.int64)])
                    assert np.array_equal(gpay[m].astype(np.int64), bpay[gb[m].astype(np.int64)])
                    assert np.array_equal(gpk.astype(np.int64), pk[gp.astype(np.int64)])
            r.release()
            r = j.probe_finish(0, want=flags | pkg.OUT_HOST)
            fb, fp = r.numpy()
            r.release()
        o1, o2 = np.lexsort((fb, fp)), np.lexsort(want_fin)
        assert np.array_equal(fb[o1], want_fin[0][o2])


@pytest.mark.parametrize("join_type", ["inner", "right", "full", "right_anti"])
@pytest.mark.parametrize("unique,layout,radix_bits", [(True, "radix", 5), (False, "radix", 3), (True, "direct", 0)])
def test_probe_pieces_equals_one_batch(pkg, join_type, unique, layout, radix_bits):
    """b200j_probe_pieces: a probe batch handed over as ragged device-resident pieces (what the multi-GPU "pull" exchange
    hands to the local join: one region per source rank) gives the pairs of the concatenated batch -- through the
    block-local partition (unique build keys behind the radix layout) and through the concatenating fall-back."""
    import torch
    rng = np.random.default_rng(hash((join_type, unique, layout)) & 0xFFFFFFFF)
    nb = 40_000
    bk = (rng.permutation(nb * 2)[:nb] if unique else rng.integers(0, nb // 4, nb)).astype(np.int64)
    lens = [70_001, 0, 33, 4096 * 8 * 3 + 5, 1, 50_000]
    pk = rng.integers(-50, nb * 2 + 50, sum(lens)).astype(np.int64)
    oj = oracle.OracleJoin(join_type, [bk])
    want = oj.probe_batch([pk], len(pk))
    dev = torch.from_numpy(np.concatenate([pk[:1], pk])).cuda()        # pieces start at odd offsets of one allocation: misaligned as well
    pieces, at = [], 1
    for m in lens:
        pieces.append((dev[at:at + m].data_ptr() if m else 0, m))
        at += m
    with pkg.HashJoin(join_type, key_widths=(8,), layout=layout, radix_bits=radix_bits) as j:
        j.build_push(0, [bk])
        j.build_finish(0)
        r = j.probe_pieces(0, pieces, want=pkg.OUT_PAIRS | pkg.OUT_HOST)
        got = r.numpy()
        r.release()
        r = j.probe_pieces(0, pieces, want=pkg.OUT_PAIRS | pkg.OUT_COUNT_ONLY)
        assert r.n == len(want[0])
        r.release()
        st = j.stats()
    o1, o2 = np.lexsort(got), np.lexsort(want)
    assert got[0].shape == want[0].shape
    assert np.array_equal(got[0][o1], want[0][o2]) and np.array_equal(got[1][o1], want[1][o2])
    if unique and layout == "radix" and join_type in ("inner", "right", "full"):
        assert st["seg_probe_launches"] >= 1


@pytest.mark.parametrize("join_type", ["inner", "full"])
def test_narrow_key_stream_falls_back_on_wide_keys(pkg, join_type):
    """Behind the radix layout 8-byte probe keys travel through the partitioned stream as 32-bit values while they fit; a batch
    with a key outside [0, 2^32) (negative, or large) must be repeated wide -- results are those of the oracle either way."""
    rng = np.random.default_rng(31)
    small = rng.permutation(200_000)[:60_000].astype(np.int64)
    wide = np.array([-1, -2**62, 2**32, 2**32 + 5, 2**63 - 1, 2**40 + 123], dtype=np.int64)
    bk = np.concatenate([small, wide])
    batches = [rng.integers(0, 220_000, 100_000).astype(np.int64),                                  # fits 32 bits
               np.concatenate([rng.integers(0, 220_000, 50_000), wide, [-5, 2**33]]).astype(np.int64),     # does not
               rng.integers(0, 220_000, 70_001).astype(np.int64)]
    oj = oracle.OracleJoin(join_type, [bk])
    with pkg.HashJoin(join_type, key_widths=(8,), layout="radix", radix_bits=4) as j:
        j.build_push(0, [bk])
        j.build_finish(0)
        for i, pb in enumerate(batches):
            r = j.probe_batch(0, [pb])
            got, want = r.numpy(), oj.probe_batch([pb], len(pb))
            r.release()
            o1, o2 = np.lexsort(got), np.lexsort(want)
            assert got[0].shape == want[0].shape, i
            assert np.array_equal(got[0][o1], want[0][o2]) and np.array_equal(got[1][o1], want[1][o2]), i
        assert j.stats()["seg_probe_launches"] == 3


@pytest.mark.parametrize("unique,layout,radix_bits", [(True, "radix", 5), (False, "radix", 3), (True, "direct", 0)])
def test_probe_pieces_with_32_bit_keys_of_a_64_bit_join(pkg, unique, layout, radix_bits):
    """piece_key_width = 4: the pieces hold the zero-extended 32-bit values of 8-byte keys (what the "pull" exchange sends
    over NVLink while every key fits 32 bits)."""
    import torch
    rng = np.random.default_rng(hash((unique, layout, "n32")) & 0xFFFFFFFF)
    nb = 50_000
    bk = (rng.permutation(nb * 2)[:nb] if unique else rng.integers(0, nb // 4, nb)).astype(np.int64)
    bk[:3] = [2**32 + 7, -9, 2**40]                                   # build keys no 32-bit probe value can reach
    lens = [40_001, 7, 4096 * 8 * 2 + 11]
    pk32 = rng.integers(0, nb * 2 + 50, sum(lens)).astype(np.uint32)
    pk32[::5000] = 0xFFFFFFF0                                          # large unsigned values stay positive when widened
    want = oracle.OracleJoin("inner", [bk]).probe_batch([pk32.astype(np.int64)], len(pk32))
    dev = torch.from_numpy(pk32.view(np.int32)).cuda()
    pieces, at = [], 0
    for m in lens:
        pieces.append((dev[at:at + m].data_ptr(), m))
        at += m
    with pkg.HashJoin("inner", key_widths=(8,), layout=layout, radix_bits=radix_bits) as j:
        j.build_push(0, [bk])
        j.build_finish(0)
        r = j.probe_pieces(0, pieces, want=pkg.OUT_PAIRS | pkg.OUT_HOST, key_width=4)
        got = r.numpy()
        r.release()
    o1, o2 = np.lexsort(got), np.lexsort(want)
    assert got[0].shape == want[0].shape
    assert np.array_equal(got[0][o1], want[0][o2]) and np.array_equal(got[1][o1], want[1][o2])


def test_guarded_buffers_stay_intact(pkg, monkeypatch):
    """compute-sanitizer is closed on this GPU pool, so the library brings its own out-of-bounds check: with B200J_GUARD=1
    every device block it allocates sits between two canaries that are verified when the block is freed; b200j_synchronize
    fails if any was overwritten. Run over every kernel family: block-local partition + segment probe (NULL keys, misaligned
    batches, 32-bit and 64-bit streams), two-pass partition + staged probe (duplicates), direct layout, composite and Utf8
    keys, the gather (plain, hinted: slot-ordered copies), probe-order results, the pieces entry point, the pipelined host path,
    the single-pass exchange scatter."""
    import torch
    from datafusion_parallelism_b200.exchange import shuffle_local_scatter, shuffle_count, shuffle_scatter
    monkeypatch.setenv("B200J_GUARD", "1")
    monkeypatch.setenv("B200J_HOST_CHUNK_ROWS", "50021")
    rng = np.random.default_rng(77)
    nb = 120_000
    uniq = rng.permutation(nb * 2)[:nb].astype(np.int64)
    dup = rng.integers(0, nb // 5, nb).astype(np.int64)
    pk = rng.integers(-10, nb * 2 + 10, 400_003).astype(np.int64)
    pk_wide = pk.copy(); pk_wide[5] = 2**41
    valid = rng.random(len(pk)) > 0.03
    for jt in ("inner", "full", "left_anti"):
        for bk, layout, bits in ((uniq, "radix", 5), (dup, "radix", 3), (uniq, "direct", 0), (dup, "direct", 0)):
            with pkg.HashJoin(jt, key_widths=(8,), build_payload_types=(8,), layout=layout, radix_bits=bits) as j:
                j.build_push(0, [bk], [bk * 3])
                j.build_finish(0)
                oj = oracle.OracleJoin(jt, [bk])
                for probe in (pa.array(pk, mask=~valid), pa.array(pk_wide).slice(3), pa.array(pk[:1])):
                    r = j.probe_batch(0, [probe], want=pkg.OUT_PAIRS | pkg.OUT_HOST)
                    got, want = r.numpy(), oj.probe_batch([probe], len(probe))
                    r.release()
                    assert got[0].shape == want[0].shape
                dpk = torch.from_numpy(np.concatenate([pk[:1], pk])).cuda()[1:]
                r = j.probe_batch(0, [dpk], want=pkg.OUT_PAIRS | pkg.OUT_BUILD_U32)
                if jt != "left_anti":
                    g = j.gather(r, [dpk], want=0)
                    g.release()
                r.release()
                if jt != "left_anti":
                    # the gather hint (key stream + slot array + slot-ordered payload copies) and the probe-order result, from
                    # device and from pipelined host batches
                    r = j.probe_batch(0, [dpk], want=pkg.OUT_PAIRS | pkg.OUT_GATHER_HINT)
                    g = j.gather(r, [dpk], want=0)
                    g.release()
                    r.release()
                    for keys, extra in (([dpk], 0), ([pa.array(pk, mask=~valid)], pkg.OUT_HOST)):
                        r = j.probe_batch(0, keys, want=pkg.OUT_PAIRS | pkg.OUT_PROBE_ORDER | pkg.OUT_BUILD_U32 | extra)
                        assert r.probe_order == (bk is uniq)
                        r.release()
                half = len(pk) // 2
                r = j.probe_pieces(0, [(dpk[:half].data_ptr(), half), (dpk[half:].data_ptr(), len(pk) - half)], want=pkg.OUT_PAIRS)
                r.release()
                j.probe_finish(0).release()
                j.synchronize()                                  # raises if a canary was overwritten
    with pkg.HashJoin("inner", key_widths=(8, 4)) as j:         # composite keys: row-hash table + equal_rows
        a8, a4 = rng.integers(0, 300, 50_000).astype(np.int64), rng.integers(0, 50, 50_000).astype(np.int32)
        j.build_push(0, [a8, a4])
        j.build_finish(0)
        j.probe_batch(0, [a8[::-1].copy(), a4[::-1].copy()]).release()
        j.synchronize()
    with pkg.HashJoin("inner", key_widths=(8,)) as j:           # exchange kernels
        dev = torch.from_numpy(pk).cuda()
        c = shuffle_count(j, dev, 3)
        bufs = [torch.empty(len(pk), dtype=torch.int64, device="cuda") for _ in range(4)]
        rids = [torch.empty(len(pk), dtype=torch.int32, device="cuda") for _ in range(4)]
        shuffle_scatter(j, 3, [t.data_ptr() for t in bufs], [t.data_ptr() for t in rids], [0] * 4)
        cap = int(c[:3].max()) + 64
        sk, sr = torch.empty(5 * cap, dtype=torch.int64, device="cuda"), torch.empty(5 * cap, dtype=torch.int32, device="cuda")
        assert np.array_equal(shuffle_local_scatter(j, dev, 3, sk.data_ptr(), sr.data_ptr(), cap)[:3], c[:3])
        j.synchronize()


@pytest.mark.parametrize("join_type", ["inner", "full", "right_anti", "left_semi"])
@pytest.mark.parametrize("layout,radix_bits", [("direct", 0), ("radix", 3)])
def test_coalescing_feeder_matches_per_batch_probing(pkg, join_type, layout, radix_bits):
    """b200j_probe_push / flush / gather_staged: DataFusion-sized probe batches (8192 rows, NULL keys, an
    Int64 and a Utf8 payload column, host memory) are staged on the device and probed in one go; the gathered rows of two
    flush rounds + the finaliser are the oracle's whole-join rows."""
    rng = np.random.default_rng(hash((join_type, layout, "feed")) & 0xFFFFFFFF)
    nb, npr = 9_000, 60_000
    build = pa.table({"id": pa.array(rng.integers(0, 6_000, nb).astype(np.int32), mask=rng.random(nb) < 0.03),
                      "value": pa.array([f"b{i % 97}" for i in range(nb)])})
    probe = pa.table({"id": pa.array(rng.integers(-100, 7_000, npr).astype(np.int32), mask=rng.random(npr) < 0.05),
                      "v": pa.array(rng.integers(0, 1 << 50, npr).astype(np.int64), mask=rng.random(npr) < 0.1),
                      "s": pa.array([None if i % 13 == 0 else "p" * (i % 9) + str(i) for i in range(npr)])})
    batches = [probe.slice(lo, 8192) for lo in range(0, npr, 8192)]
    want = oracle.join_tables(build, batches, [("id", "id")], join_type)
    left_only, right_only = join_type in ("left_semi", "left_anti"), join_type in ("right_semi", "right_anti")
    types = ([] if right_only else [pa.int32(), pa.string()]) + ([] if left_only else [pa.int32(), pa.int64(), pa.string()])
    pieces = []
    with pkg.HashJoin(join_type, key_widths=(4,), build_payload_types=(pkg.T_UTF8,), probe_payload_types=(8, pkg.T_UTF8),
                      layout=layout, radix_bits=radix_bits) as j:
        j.build_push(0, [build.column("id")], [build.column("value")])
        j.build_finish(0)
        for rnd in (batches[:5], batches[5:]):
            for pb in rnd:
                pb = pb.combine_chunks()
                j.probe_push(0, [pb.column("id")], [pb.column("v"), pb.column("s")])
            assert j.staged_rows(0) == sum(b.num_rows for b in rnd)
            res = j.probe_flush(0, want=pkg.OUT_PAIRS)
            g = j.gather_staged(0, res)
            assert j.staged_rows(0) == 0
            if not left_only:
                pieces.append(pa.Table.from_arrays(g.to_arrow(types), names=want.column_names))
            g.release()
            res.release()
        fin = j.probe_finish(0, want=pkg.OUT_PAIRS | pkg.OUT_HOST)
        if fin.n or not pieces:
            g = j.gather(fin, None)
            pieces.append(pa.Table.from_arrays(g.to_arrow(types), names=want.column_names))
            g.release()
        fin.release()
    got = pa.concat_tables(pieces)
    assert got.num_rows == want.num_rows
    assert rows_multiset(got) == rows_multiset(want)
