"""b200j_shuffle_count_static: the allocation-free variant of b200j_shuffle_count (scratch in a static device array of the
library; no counterpart in the reference). Collected last on purpose: it must pass alone in
a fresh process and after every other test. Called through the C ABI; the result must be IDENTICAL (integer counts) to the
allocating entry point's, on calls in a row that share the static array."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ERR_INVALID, ERR_UNSUPPORTED, ERR_STATE = -1, -4, -5          # include/b200join.h


@pytest.fixture(scope="module")
def env():
    import torch
    import datafusion_parallelism_b200 as d
    assert d.device_count() >= 1, "no CUDA device: -m gpu tests must run on the B200 box"
    return d, torch


def _static_counts(d, torch, j, keys_t, n_rows, n_dest, validity_t=None, bit_offset=0, counts_t=None):
    """One call of the new entry point straight on the ctypes symbol; returns the counts as numpy (read after synchronize)."""
    from datafusion_parallelism_b200 import _native as N
    if counts_t is None:
        counts_t = torch.full((n_dest + 1,), -7, dtype=torch.int64, device="cuda")     # garbage: the call must overwrite all of it
    torch.cuda.synchronize()
    rc = N.lib().b200j_shuffle_count_static(j._h, C.c_void_p(keys_t.data_ptr()) if n_rows else None,
                                            C.c_void_p(validity_t.data_ptr()) if validity_t is not None else None,
                                            bit_offset, n_rows, n_dest, C.c_void_p(counts_t.data_ptr()))
    assert rc == 0, N.lib().b200j_last_error(j._h)
    j.synchronize()
    return counts_t.cpu().numpy().copy()


@pytest.mark.parametrize("dtype", [np.int64, np.int32])
def test_static_count_matches_allocating_count_twice_in_a_row(env, dtype):
    d, torch = env
    from datafusion_parallelism_b200.exchange import shuffle_count
    rng = np.random.default_rng(11)
    kw = np.dtype(dtype).itemsize
    with d.HashJoin("inner", key_widths=(kw,)) as j:
        # two calls in a row with DIFFERENT sizes and destination counts: the second call finds the first one's histograms
        # in the static array (more columns than it uses itself) and must not see them
        cases = [(1_000_003, 8), (70_001, 3), (1_000_003, 8), (5, 16), (300_000, 1)]
        for n, n_dest in cases:
            keys = rng.integers(-(1 << 40) if kw == 8 else -(1 << 30), (1 << 40) if kw == 8 else (1 << 30), n).astype(dtype)
            dev = torch.from_numpy(keys).cuda()
            want = shuffle_count(j, dev, n_dest)[: n_dest + 1]
            got1 = _static_counts(d, torch, j, dev, n, n_dest)
            got2 = _static_counts(d, torch, j, dev, n, n_dest)
            assert got1.sum() == n
            assert np.array_equal(got1, want), (n, n_dest, got1, want)
            assert np.array_equal(got2, want)


def test_static_count_nulls_unaligned_and_empty(env):
    d, torch = env
    from datafusion_parallelism_b200.exchange import shuffle_count
    from datafusion_parallelism_b200.join import ColumnRef
    rng = np.random.default_rng(12)
    n, n_dest, bit_offset = 200_001, 5, 3
    keys = rng.integers(0, 1 << 50, n + 1).astype(np.int64)
    valid = rng.random(n) < 0.8
    bits = np.zeros(bit_offset + n, dtype=bool)
    bits[bit_offset:] = valid
    bitmap = np.packbits(bits, bitorder="little")
    with d.HashJoin("inner", key_widths=(8,)) as j:
        dev_all = torch.from_numpy(keys).cuda()
        dev = dev_all[1:]                                   # 8 bytes off the 16-byte boundary: the kernel's "mis" path
        vdev = torch.from_numpy(bitmap).cuda()
        col = ColumnRef(dev.data_ptr(), None, vdev.data_ptr(), bit_offset, n, d.MEM_DEVICE, 8, (dev_all, vdev))
        want = shuffle_count(j, col, n_dest)[: n_dest + 1]
        assert want[n_dest] == int((~valid).sum())
        for _ in range(2):
            got = _static_counts(d, torch, j, dev, n, n_dest, validity_t=vdev, bit_offset=bit_offset)
            assert np.array_equal(got, want)
        # no rows: zeros, no kernel reads the (NULL) key pointer
        got = _static_counts(d, torch, j, dev, 0, n_dest)
        assert np.array_equal(got, np.zeros(n_dest + 1, dtype=np.int64))
        # and the allocating path still works after the static one
        assert np.array_equal(shuffle_count(j, col, n_dest)[: n_dest + 1], want)


def test_static_count_rejects_what_it_cannot_hold(env):
    d, torch = env
    from datafusion_parallelism_b200 import _native as N
    from datafusion_parallelism_b200.exchange import STATIC_MAX_ROWS, shuffle_set_hot
    lib = N.lib()
    with d.HashJoin("inner", key_widths=(8,)) as j:
        dev = torch.zeros(16, dtype=torch.int64, device="cuda")
        counts = torch.zeros(17, dtype=torch.int64, device="cuda")
        kp, cp = C.c_void_p(dev.data_ptr()), C.c_void_p(counts.data_ptr())
        # validated before anything is launched: the 16-row buffer is never read
        assert lib.b200j_shuffle_count_static(j._h, kp, None, 0, STATIC_MAX_ROWS + 1, 4, cp) == ERR_INVALID
        assert b"B200J_STATIC_MAX_ROWS" in lib.b200j_last_error(j._h)
        assert lib.b200j_shuffle_count_static(j._h, kp, None, 0, -1, 4, cp) == ERR_INVALID
        assert lib.b200j_shuffle_count_static(j._h, kp, None, 0, 16, 0, cp) == ERR_INVALID
        assert lib.b200j_shuffle_count_static(j._h, kp, None, 0, 16, 17, cp) == ERR_INVALID
        assert lib.b200j_shuffle_count_static(j._h, kp, None, 0, 16, 4, None) == ERR_INVALID
        assert lib.b200j_shuffle_count_static(j._h, None, None, 0, 16, 4, cp) == ERR_INVALID
        assert lib.b200j_shuffle_count_static(j._h, C.c_void_p(dev.data_ptr() + 4), None, 0, 8, 4, cp) == ERR_INVALID
        shuffle_set_hot(j, [1, 2, 3])
        assert lib.b200j_shuffle_count_static(j._h, kp, None, 0, 16, 4, cp) == ERR_STATE
        shuffle_set_hot(j, None)
        assert lib.b200j_shuffle_count_static(j._h, kp, None, 0, 16, 4, cp) == 0
        j.synchronize()
        assert int(counts[:5].sum().item()) == 16
    with d.HashJoin("inner", key_widths=(d.T_UTF8,)) as j:
        assert lib.b200j_shuffle_count_static(j._h, kp, None, 0, 16, 4, cp) == ERR_UNSUPPORTED


def test_static_count_at_the_maximum(env):
    """B200J_STATIC_MAX_ROWS rows from a buffer that starts 8 bytes off a 16-byte boundary: the largest grid the array holds."""
    d, torch = env
    from datafusion_parallelism_b200.exchange import STATIC_MAX_ROWS, shuffle_count, _dev_column
    n, n_dest = STATIC_MAX_ROWS, 16
    g = torch.Generator(device="cuda").manual_seed(3)
    dev_all = torch.randint(-(1 << 62), 1 << 62, (n + 1,), dtype=torch.int64, device="cuda", generator=g)
    dev = dev_all[1:]
    with d.HashJoin("inner", key_widths=(8,)) as j:
        want = shuffle_count(j, _dev_column(dev.data_ptr(), n, 8, keep=dev_all), n_dest)[: n_dest + 1]
        for _ in range(2):
            got = _static_counts(d, torch, j, dev, n, n_dest)
            assert got.sum() == n and np.array_equal(got, want)
