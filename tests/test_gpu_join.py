"""GPU parity tests: the CUDA path, called through the C ABI, against the CPU oracle on the same seeded inputs.
Bit-exact: index pairs are compared as order-normalised multisets, hashes value by value."""
import numpy as np
import pyarrow as pa
import pytest

import oracle

pytestmark = pytest.mark.gpu

JOIN_TYPES = ["inner", "left", "right", "full", "left_semi", "right_semi", "left_anti", "right_anti"]
LAYOUT_CASES = [("direct", 0), ("radix", 3), ("radix", 7)]


@pytest.fixture(scope="module")
def pkg():
    import datafusion_parallelism_b200 as d
    assert d.device_count() >= 1, "no CUDA device: -m gpu tests must run on the B200 box"
    return d


def norm(b, p):
    b = np.asarray(b, dtype=np.uint64)
    p = np.asarray(p, dtype=np.uint32)
    o = np.lexsort((b, p))
    return b[o], p[o]


def assert_pairs_equal(got, want, ctx=""):
    gb, gp = norm(*got)
    wb, wp = norm(*want)
    assert gb.shape == wb.shape, f"{ctx}: {gb.shape[0]} rows, oracle {wb.shape[0]}"
    assert np.array_equal(gb, wb) and np.array_equal(gp, wp), ctx


def np_validity(mask_valid):
    return np.packbits(np.asarray(mask_valid, dtype=np.uint8), bitorder="little")


def as_arrow(values, valid, typ):
    return pa.array(values, type=typ, mask=None if valid is None else ~np.asarray(valid, dtype=bool))


def run_gpu(pkg, join_type, build_batches, probe_batches, key_width, layout, radix_bits, n_partitions=1):
    """build_batches / probe_batches: lists of pyarrow arrays (nulls allowed). Returns per-probe-batch pairs + finaliser."""
    out = []
    with pkg.HashJoin(join_type, key_widths=(key_width,), layout=layout, radix_bits=radix_bits, n_partitions=n_partitions) as j:
        for i, bb in enumerate(build_batches):
            j.build_push(i % n_partitions, [bb])
        for part in range(n_partitions):
            j.build_finish(part)
        for i, pb in enumerate(probe_batches):
            r = j.probe_batch(i % n_partitions, [pb])
            out.append(r.numpy())
            r.release()
        fins = []
        for part in range(n_partitions):
            r = j.probe_finish(part)
            fins.append(r.numpy())
            r.release()
        stats = j.stats()
    return out, fins, stats


def run_oracle(join_type, build_all, probe_batches):
    oj = oracle.OracleJoin(join_type, [build_all])
    out = [oj.probe_batch([pb], len(pb)) for pb in probe_batches]
    return out, oj.finish()


def gen_case(rng, n_build, n_probe, key_range, null_frac, dtype):
    bk = rng.integers(0, key_range, n_build).astype(dtype)
    pk = rng.integers(-key_range // 8, key_range + key_range // 8, n_probe).astype(dtype)
    bv = rng.random(n_build) >= null_frac if null_frac > 0 else None
    pv = rng.random(n_probe) >= null_frac if null_frac > 0 else None
    return bk, bv, pk, pv


@pytest.mark.parametrize("layout,radix_bits", LAYOUT_CASES)
@pytest.mark.parametrize("join_type", JOIN_TYPES)
@pytest.mark.parametrize("dtype,null_frac", [(np.int64, 0.0), (np.int64, 0.1), (np.int32, 0.05)])
def test_pairs_match_oracle(pkg, join_type, layout, radix_bits, dtype, null_frac):
    rng = np.random.default_rng(hash((join_type, layout, radix_bits, str(dtype), null_frac)) & 0xFFFFFFFF)
    typ = pa.int64() if dtype == np.int64 else pa.int32()
    bk, bv, pk, pv = gen_case(rng, 20_000, 50_000, 15_000, null_frac, dtype)   # duplicates on both sides, key 0 present
    build_batches = [as_arrow(bk[i:i + 7000], None if bv is None else bv[i:i + 7000], typ) for i in range(0, len(bk), 7000)]
    probe_batches = [as_arrow(pk[i:i + 16_384], None if pv is None else pv[i:i + 16_384], typ) for i in range(0, len(pk), 16_384)]
    got, fins, stats = run_gpu(pkg, join_type, build_batches, probe_batches, np.dtype(dtype).itemsize, layout, radix_bits)
    want, want_fin = run_oracle(join_type, as_arrow(bk, bv, typ), probe_batches)
    for i, (g, w) in enumerate(zip(got, want)):
        assert_pairs_equal(g, w, f"{join_type} batch {i}")
    assert_pairs_equal(fins[-1], want_fin, f"{join_type} finaliser")
    assert stats["build_rows"] == len(bk)
    assert stats["layout"] == (1 if layout == "direct" else 2)


@pytest.mark.parametrize("layout,radix_bits", LAYOUT_CASES)
@pytest.mark.parametrize("join_type", JOIN_TYPES)
def test_pipelined_host_probe(pkg, monkeypatch, join_type, layout, radix_bits):
    """Host batch in, host pairs out: the library cuts the batch into chunks and overlaps H2D / kernels / D2H
    (probe_host_pipelined). Chunk size shrunk so one batch is ~17 ragged chunks; NULLs with a non-zero bitmap offset;
    duplicate build keys so the pinned result has to grow; the result must be that of the unchunked call."""
    monkeypatch.setenv("B200J_HOST_CHUNK_ROWS", "3001")
    rng = np.random.default_rng(hash((join_type, layout, "pipe")) & 0xFFFFFFFF)
    bk, bv, pk, pv = gen_case(rng, 20_000, 50_003, 6_000, 0.07, np.int64)
    probe = as_arrow(pk, pv, pa.int64()).slice(5)               # Arrow array offset 5: bitmap bit offset, unaligned keys
    got, fins, stats = run_gpu(pkg, join_type, [as_arrow(bk, bv, pa.int64())], [probe], 8, layout, radix_bits)
    want, want_fin = run_oracle(join_type, as_arrow(bk, bv, pa.int64()), [probe])
    assert_pairs_equal(got[0], want[0], f"{join_type} pipelined")
    assert_pairs_equal(fins[-1], want_fin, f"{join_type} finaliser")
    with pkg.HashJoin(join_type, key_widths=(8,), layout=layout, radix_bits=radix_bits) as j:
        j.build_push(0, [as_arrow(bk, bv, pa.int64())])
        j.build_finish(0)
        r = j.probe_batch(0, [probe], want=pkg.OUT_PAIRS | pkg.OUT_COUNT_ONLY)
        assert r.n == len(want[0][1])
        r.release()


@pytest.mark.parametrize("layout,radix_bits", [("direct", 0), ("radix", 5)])
@pytest.mark.parametrize("join_type", JOIN_TYPES)
@pytest.mark.parametrize("collide", [0, 2])
def test_composite_keys(pkg, join_type, layout, radix_bits, collide):
    """Two key columns (Int64, Int32) with NULLs in either (TPC-H Q9's (ps_partkey, ps_suppkey) shape): the table is keyed on
    the combined row hash, equality is checked on the candidate pairs. collide=2 truncates the row hash to 10 bits so that
    almost every candidate is a hash collision the equality filter has to drop."""
    rng = np.random.default_rng(hash((join_type, layout, collide, "comp")) & 0xFFFFFFFF)
    nb, npr = 12_000, 30_000
    b0, b1 = rng.integers(0, 300, nb).astype(np.int64), rng.integers(0, 40, nb).astype(np.int32)
    p0, p1 = rng.integers(-20, 330, npr).astype(np.int64), rng.integers(0, 44, npr).astype(np.int32)
    bv0, bv1 = rng.random(nb) >= 0.05, rng.random(nb) >= 0.05
    pv0, pv1 = rng.random(npr) >= 0.05, rng.random(npr) >= 0.05
    build = [as_arrow(b0, bv0, pa.int64()), as_arrow(b1, bv1, pa.int32())]
    probes = [[as_arrow(p0[lo:lo + 11_000], pv0[lo:lo + 11_000], pa.int64()), as_arrow(p1[lo:lo + 11_000], pv1[lo:lo + 11_000], pa.int32())]
              for lo in range(0, npr, 11_000)]
    oj = oracle.OracleJoin(join_type, build)
    with pkg.HashJoin(join_type, key_widths=(8, 4), layout=layout, radix_bits=radix_bits, hash_flags=collide) as j:
        j.build_push(0, build)
        j.build_finish(0)
        for i, pb in enumerate(probes):
            r = j.probe_batch(0, pb)
            assert_pairs_equal(r.numpy(), oj.probe_batch(pb, len(pb[0])), f"{join_type} batch {i}")
            r.release()
        r = j.probe_finish(0)
        assert_pairs_equal(r.numpy(), oj.finish(), f"{join_type} finaliser")
        r.release()
        if collide == 0:
            assert np.array_equal(j.hash_keys(probes[0]), oracle.create_hashes(probes[0]))


UTF8_WORDS = ["", "a", "ab", "abc", "abcd", "abcde", "abcdefg", "abcdefgh", "abcdefghi", "0123456789abcdef", "0123456789abcdefg",
              "x" * 31, "y" * 32, "z" * 33, "naïve", "日本語のキー"] + [f"customer#{i:09d}" for i in range(400)]


def utf8_array(rng, n, null_rate=0.05, hi=None):
    idx = rng.integers(0, hi or len(UTF8_WORDS), n)
    return pa.array([UTF8_WORDS[i] for i in idx], type=pa.string(), mask=rng.random(n) < null_rate)


@pytest.mark.parametrize("layout,radix_bits", [("direct", 0), ("radix", 5)])
@pytest.mark.parametrize("join_type", JOIN_TYPES)
@pytest.mark.parametrize("collide", [0, 2])
def test_utf8_keys(pkg, join_type, layout, radix_bits, collide):
    """Utf8 join key ("hash on device, equality via offsets"): key_widths = (B200J_T_UTF8,). The table is
    keyed on the string's row hash, equal_rows compares the bytes; duplicates and NULLs on both sides, probe words the build
    side does not have, sliced (offset != 0) probe batches."""
    rng = np.random.default_rng(hash((join_type, layout, collide, "utf8")) & 0xFFFFFFFF)
    nb, npr = 6_000, 20_000
    build = [utf8_array(rng, nb, hi=300)]
    whole = utf8_array(rng, npr)
    probes = [[whole.slice(lo, 7_000)] for lo in range(0, npr, 7_000)]
    oj = oracle.OracleJoin(join_type, build)
    with pkg.HashJoin(join_type, key_widths=(pkg.T_UTF8,), layout=layout, radix_bits=radix_bits, hash_flags=collide) as j:
        j.build_push(0, [build[0].slice(0, 2_500)])
        j.build_push(0, [build[0].slice(2_500)])
        j.build_finish(0)
        for i, pb in enumerate(probes):
            r = j.probe_batch(0, pb)
            assert_pairs_equal(r.numpy(), oj.probe_batch(pb, len(pb[0])), f"{join_type} batch {i}")
            r.release()
        r = j.probe_finish(0)
        assert_pairs_equal(r.numpy(), oj.finish(), f"{join_type} finaliser")
        r.release()
        if collide == 0:
            for pb in probes[:2]:
                assert np.array_equal(j.hash_keys(pb), oracle.create_hashes(pb))


def test_utf8_with_int_key_device_batches(pkg):
    """(Utf8, Int32) composite key from DEVICE buffers: data bytes, int32 offsets and the bitmap live in HBM."""
    import torch
    rng = np.random.default_rng(77)
    nb, npr = 3_000, 9_000

    def side(n):
        s = utf8_array(rng, n, hi=120)
        k2 = pa.array(rng.integers(0, 4, n).astype(np.int32), mask=rng.random(n) < 0.03)
        return s, k2
    bs, b2 = side(nb)
    ps, p2 = side(npr)
    want = oracle.OracleJoin("inner", [bs, b2])

    def dev_cols(s, k2):
        bufs = s.buffers()
        n = len(s)
        off = torch.from_numpy(np.frombuffer(bufs[1], dtype=np.int32)[: n + 1].copy()).cuda()
        data = torch.from_numpy(np.frombuffer(bufs[2], dtype=np.uint8).copy()).cuda()
        val = torch.from_numpy(np.packbits(np.asarray(s.is_valid()).astype(np.uint8), bitorder="little")).cuda()
        c0 = pkg.ColumnRef(data.data_ptr(), off.data_ptr(), val.data_ptr(), 0, n, pkg.MEM_DEVICE, pkg.T_UTF8, (off, data, val))
        v2 = torch.from_numpy(np.packbits(np.asarray(k2.is_valid()).astype(np.uint8), bitorder="little")).cuda()
        c1 = pkg.column(torch.from_numpy(k2.fill_null(0).to_numpy(zero_copy_only=False).astype(np.int32)).cuda(), validity=v2)
        return [c0, c1]
    with pkg.HashJoin("inner", key_widths=(pkg.T_UTF8, 4), layout="direct") as j:
        j.build_push(0, dev_cols(bs, b2))
        j.build_finish(0)
        r = j.probe_batch(0, dev_cols(ps, p2), want=pkg.OUT_PAIRS | pkg.OUT_HOST)
        assert_pairs_equal(r.numpy(), want.probe_batch([ps, p2], npr), "utf8+int32 device")
        r.release()


@pytest.mark.parametrize("join_type", ["inner", "left", "right", "full"])
@pytest.mark.parametrize("dtype", [np.int64, np.int32])
@pytest.mark.parametrize("n_build", [40, 60_000])
def test_dense_direct_paths(pkg, join_type, dtype, n_build):
    """Unique build keys on the direct layout from DEVICE buffers: the un-compacted dense kernel (all hits), dense + compaction
    (first batch with misses), the staged kernel once the previous batch had > 30 % misses, a misaligned key slice (staged
    kernel again: the dense one reads 16 B vectors), and back to dense. n_build = 40 also takes the L1-allocating loads of
    tiny tables. Every batch is compared with the oracle."""
    import torch
    rng = np.random.default_rng(hash((join_type, str(dtype), n_build, "dd")) & 0xFFFFFFFF)
    bk = (rng.permutation(n_build * 3)[:n_build]).astype(dtype)              # unique, key 0 may be present
    hits = lambda n: bk[rng.integers(0, n_build, n)]
    mostly_misses = lambda n: np.where(rng.random(n) < 0.1, bk[rng.integers(0, n_build, n)], rng.integers(n_build * 3, n_build * 6, n)).astype(dtype)
    batches = [hits(30_001), mostly_misses(30_000), mostly_misses(25_003), hits(20_002), hits(20_000), hits(7)]
    oj = oracle.OracleJoin(join_type, [bk])
    with pkg.HashJoin(join_type, key_widths=(bk.dtype.itemsize,), layout="direct") as j:
        j.build_push(0, [torch.from_numpy(bk).cuda()])
        j.build_finish(0)
        launches = []
        for i, pb in enumerate(batches):
            t = torch.from_numpy(np.concatenate([pb[:1], pb])).cuda()
            dev = t[1:] if i == 3 else t[1:].clone()                          # batch 3: pointer 4 / 8 B past a 16 B boundary
            r = j.probe_batch(0, [dev], want=pkg.OUT_PAIRS | pkg.OUT_HOST)
            assert_pairs_equal(r.numpy(), oj.probe_batch([pb], len(pb)), f"{join_type} batch {i}")
            r.release()
            launches.append(j.stats()["dense_probe_launches"])
        r = j.probe_finish(0)
        assert_pairs_equal(r.numpy(), oj.finish(), f"{join_type} finaliser")
        r.release()
    dense = [b - a for a, b in zip([0] + launches[:-1], launches)]
    if join_type in ("inner", "left"):
        assert dense == [1, 1, 0, 0, 1, 1], dense         # batch 2: staged after 90 % misses; batch 3: misaligned keys
    else:
        assert dense == [1, 1, 1, 0, 1, 1], dense         # Right / Full keep the NULL rows: dense at any miss rate


@pytest.mark.parametrize("layout,radix_bits", LAYOUT_CASES)
def test_unique_keys_one_to_one(pkg, layout, radix_bits):
    """The BASELINE C1/C2 shape in miniature: unique build keys (a permutation), uniform probe, every probe row matches once."""
    n_build, n_probe = 1 << 18, 1 << 20
    bk = oracle.gen_permutation_i64(n_build, 3)
    pk = oracle.gen_uniform_i64(n_probe, n_build, 4)
    got, fins, stats = run_gpu(pkg, "inner", [bk], [pk], 8, layout, radix_bits)
    b, p = got[0]
    assert len(b) == n_probe and stats["distinct_keys"] == n_build and stats["duplicate_rows"] == 0
    assert np.array_equal(np.sort(p), np.arange(n_probe, dtype=np.uint32))
    assert np.array_equal(bk[b.astype(np.int64)], pk[p.astype(np.int64)])       # key equality on every pair
    want, _ = run_oracle("inner", bk, [pk])
    assert_pairs_equal(got[0], want[0])


@pytest.mark.parametrize("layout,radix_bits", LAYOUT_CASES)
def test_heavy_duplicate_chains(pkg, layout, radix_bits):
    """Skew: a few keys own most build rows (runs longer than a warp), so the warp-cooperative run copy is exercised."""
    rng = np.random.default_rng(11)
    heavy = np.repeat(np.array([7, 0, 123456789012], dtype=np.int64), [5000, 300, 40])
    light = rng.integers(1000, 3000, 4000).astype(np.int64)
    bk = rng.permutation(np.concatenate([heavy, light]))
    pk = np.concatenate([np.array([7, 7, 0, 123456789012, 5, 7], dtype=np.int64), rng.integers(900, 3100, 3000).astype(np.int64)])
    for jt in ("inner", "left", "full", "right_semi", "left_anti"):
        got, fins, stats = run_gpu(pkg, jt, [bk], [pk], 8, layout, radix_bits)
        want, want_fin = run_oracle(jt, bk, [pk])
        assert_pairs_equal(got[0], want[0], jt)
        assert_pairs_equal(fins[-1], want_fin, jt + " finaliser")
    assert stats["duplicate_rows"] > 5000


def test_exponential_generator_keys(pkg):
    """benches/exponential_distribution.rs AllEqualSize shape: make_exponential_int_array per 1024-row block."""
    blocks = [oracle.make_exponential_int_array(i * 1024, (i + 1) * 1024) for i in range(10)]
    bk = np.concatenate(blocks).astype(np.int32)
    pk = np.arange(0, 10 * 1024, dtype=np.int32)
    for layout, bits in LAYOUT_CASES:
        got, fins, _ = run_gpu(pkg, "inner", [bk], [pk], 4, layout, bits)
        want, _ = run_oracle("inner", bk, [pk])
        assert_pairs_equal(got[0], want[0], layout)


def test_empty_and_ragged_inputs(pkg):
    e64 = np.zeros(0, dtype=np.int64)
    some = np.array([1, 2, 3, 0, -1], dtype=np.int64)
    # empty build side
    for jt in JOIN_TYPES:
        got, fins, _ = run_gpu(pkg, jt, [e64], [some, e64], 8, "auto", 0)
        want, want_fin = run_oracle(jt, e64, [some, e64])
        for g, w in zip(got, want):
            assert_pairs_equal(g, w, jt + " empty build")
        assert_pairs_equal(fins[-1], want_fin, jt)
    # empty probe side, odd sizes, unaligned slices (arrow offset != 0 -> 8-byte-aligned-only pointers and bit offsets)
    arr = pa.array([None, 5, 6, None, 7, 5, 0, 0, 9, 11, None], type=pa.int64())
    for jt in JOIN_TYPES:
        for lo in (0, 1, 3):
            pb = arr.slice(lo)
            got, fins, _ = run_gpu(pkg, jt, [arr.slice(1, 7), arr.slice(8)], [pb, arr.slice(2, 0)], 8, "direct", 0)
            build_all = pa.concat_arrays([arr.slice(1, 7), arr.slice(8)])
            want, want_fin = run_oracle(jt, build_all, [pb, arr.slice(2, 0)])
            for g, w in zip(got, want):
                assert_pairs_equal(g, w, f"{jt} slice {lo}")
            assert_pairs_equal(fins[-1], want_fin, jt)


def test_multi_partition_protocol(pkg):
    """N partitions share one build (Arc<OnceLock<executor>>, parallel_hash_join.rs:144-151): the last build_finish
    builds, only the last probe_finish emits the finaliser, and a second finish is 'State already consumed'."""
    rng = np.random.default_rng(5)
    bk = rng.integers(0, 500, 2000).astype(np.int64)
    pks = [rng.integers(0, 700, 1000).astype(np.int64) for _ in range(6)]
    got, fins, _ = run_gpu(pkg, "full", [bk[i:i + 250] for i in range(0, 2000, 250)], pks, 8, "direct", 0, n_partitions=4)
    want, want_fin = run_oracle("full", bk, pks)
    for g, w in zip(got, want):
        assert_pairs_equal(g, w)
    assert all(len(f[0]) == 0 for f in fins[:-1])
    assert_pairs_equal(fins[-1], want_fin)
    with pkg.HashJoin("inner", n_partitions=2) as j:
        j.build_push(0, [bk])
        j.build_finish(0)
        with pytest.raises(pkg.B200JoinError) as e:
            j.build_finish(0)
        assert e.value.code == -5
        with pytest.raises(pkg.B200JoinError):
            j.probe_batch(0, [pks[0]])          # partition 1 has not finished the build side
        j.build_finish(1)
        j.probe_batch(1, [pks[0]]).release()


def test_join_filter_round_trip(pkg):
    """full_join_with_filter shape (src/lib.rs:651-726): matches out -> caller filters -> apply_pairs does the join-type algebra."""
    rng = np.random.default_rng(9)
    bk = rng.integers(0, 300, 1500).astype(np.int64)
    pk = rng.integers(0, 400, 2500).astype(np.int64)
    bval = rng.integers(0, 5, 1500)
    pval = rng.integers(0, 5, 2500)
    for jt in JOIN_TYPES:
        oj = oracle.OracleJoin(jt, [bk])
        fn = lambda b, p, n: bval[b.astype(np.int64)] != pval[p.astype(np.int64)]
        want = oj.probe_batch([pk], len(pk), fn)
        want_fin = oj.finish()
        with pkg.HashJoin(jt) as j:
            j.build_push(0, [bk])
            j.build_finish(0)
            r = j.probe_batch(0, [pk], want=pkg.OUT_PAIRS | pkg.OUT_HOST | pkg.OUT_MATCHES_ONLY)
            b, p = r.numpy()
            r.release()
            keep = fn(b, p, len(pk))
            r = j.apply_pairs(0, len(pk), b[keep], p[keep])
            got = r.numpy()
            r.release()
            r = j.probe_finish(0)
            got_fin = r.numpy()
            r.release()
        assert_pairs_equal(got, want, jt)
        assert_pairs_equal(got_fin, want_fin, jt + " finaliser")


@pytest.mark.parametrize("flags", [0, 1])
def test_hash_bit_identical_to_oracle(pkg, flags):
    """create_hashes (a1): GPU == oracle bit for bit, for 1..3 key columns, both widths, with NULLs, both ahash variants."""
    rng = np.random.default_rng(21 + flags)
    n = 100_003
    c64 = rng.integers(-2**62, 2**62, n).astype(np.int64)
    c32 = rng.integers(-2**31, 2**31 - 1, n).astype(np.int32)
    c64b = rng.integers(0, 10, n).astype(np.int64)
    v = rng.random(n) > 0.2
    a64 = as_arrow(c64, v, pa.int64())
    a32 = as_arrow(c32, ~v | (rng.random(n) > 0.5), pa.int32())
    keys = oracle.keys_with_seed(0)
    for cols, widths in (([c64], (8,)), ([a64], (8,)), ([c32], (4,)), ([a64, a32], (8, 4)), ([a32, c64b, a64], (4, 8, 8))):
        with pkg.HashJoin("inner", key_widths=widths, hash_flags=flags) as j:
            got = j.hash_keys(cols)
        want = oracle.create_hashes(cols, keys, flags)
        assert np.array_equal(got, want)
    # edge values
    edge = np.array([0, 1, -1, 2**63 - 1, -2**63, 0x243f6a8885a308d3], dtype=np.int64)
    with pkg.HashJoin("inner", hash_flags=flags) as j:
        assert np.array_equal(j.hash_keys([edge]), oracle.create_hashes([edge], keys, flags))
    # custom RandomState keys
    ck = np.array([1, 2, 3, 4], dtype=np.uint64)
    with pkg.HashJoin("inner", hash_keys=ck, hash_flags=flags) as j:
        assert np.array_equal(j.hash_keys([c64[:1000]]), oracle.create_hashes([c64[:1000]], ck, flags))


def test_device_resident_io(pkg):
    """north_star: Arrow buffers handed over as device pointers; results stay on the device."""
    import torch
    n_build, n_probe = 1 << 16, 1 << 18
    bk = torch.from_numpy(oracle.gen_permutation_i64(n_build, 1)).cuda()
    pk = torch.from_numpy(oracle.gen_uniform_i64(n_probe, n_build, 2)).cuda()
    with pkg.HashJoin("inner", timing=True) as j:
        j.build_push(0, [bk])
        j.build_finish(0)
        r = j.probe_batch(0, [pk], want=pkg.OUT_PAIRS)
        assert r.mem == pkg.MEM_DEVICE and r.n == n_probe
        b, p = r.device_arrays()
        tb = torch.as_tensor(b, device="cuda")
        tp = torch.as_tensor(p, device="cuda")
        assert torch.equal(bk[tb], pk[tp.long()])
        assert int(tp.long().sum()) == n_probe * (n_probe - 1) // 2
        r.release()
        st = j.stats()
        assert st["ms_probe"] > 0 and st["probe_launches"] >= 1
