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
