"""CPU oracle for the hash-join hot path — TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import this package.  The product (``datafusion_parallelism_b200``) never does, and
``tests/test_boundary.py`` greps for that.

The arithmetic lives in ``join_oracle.c`` / ``cpu_join_mt.c`` (each function cites the reference
file:line it restates); this module is a ctypes + numpy + pyarrow veneer over it.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Callable, List, Optional, Sequence, Tuple

import numpy as np
import pyarrow as pa

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "liborc.so")

NULL64 = np.uint64(0xFFFFFFFFFFFFFFFF)
NULL32 = np.uint32(0xFFFFFFFF)

JOIN_TYPES = {
    "inner": 0, "left": 1, "right": 2, "full": 3,
    "left_semi": 4, "right_semi": 5, "left_anti": 6, "right_anti": 7,
}


def build(force: bool = False) -> str:
    """Compile the oracle with gcc (Makefile in this directory). Returns the .so path."""
    srcs = [os.path.join(_HERE, f) for f in ("join_oracle.c", "cpu_join_mt.c", "Makefile")]
    stale = (not os.path.exists(_LIB_PATH)) or any(
        os.path.getmtime(s) > os.path.getmtime(_LIB_PATH) for s in srcs)
    if force or stale:
        subprocess.run(["make", "-C", _HERE, "-B"], check=True, stdout=subprocess.DEVNULL)
    return _LIB_PATH


class _Col(C.Structure):
    _fields_ = [("values", C.c_void_p), ("validity", C.c_void_p), ("width", C.c_int32), ("_pad", C.c_int32), ("offsets", C.c_void_p)]


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        u64p, u32p, i64p, i32p = (C.POINTER(C.c_uint64), C.POINTER(C.c_uint32),
                                  C.POINTER(C.c_int64), C.POINTER(C.c_int32))
        L.orc_keys_with_seed.argtypes = [C.c_uint64, u64p]
        L.orc_hash_u64.argtypes = [u64p, C.c_uint64, C.c_int]
        L.orc_hash_u64.restype = C.c_uint64
        L.orc_create_hashes.argtypes = [C.c_int32, C.POINTER(_Col), C.c_uint64, u64p, C.c_int32, u64p]
        L.orc_map_new.restype = C.c_void_p
        L.orc_map_free.argtypes = [C.c_void_p]
        L.orc_map_rows.argtypes = [C.c_void_p]
        L.orc_map_rows.restype = C.c_uint64
        L.orc_map_add.argtypes = [C.c_void_p, u64p, C.c_uint64]
        L.orc_map_get_all.argtypes = [C.c_void_p, C.c_uint64, u64p, C.c_uint64]
        L.orc_map_get_all.restype = C.c_uint64
        L.orc_bitset_new.argtypes = [C.c_uint64]
        L.orc_bitset_new.restype = C.c_void_p
        L.orc_bitset_free.argtypes = [C.c_void_p]
        L.orc_bitset_set_ones.argtypes = [C.c_void_p, u64p, C.c_uint64]
        L.orc_bitset_set_ones.restype = C.c_int32
        L.orc_bitset_set_count.argtypes = [C.c_void_p]
        L.orc_bitset_set_count.restype = C.c_uint64
        L.orc_bitset_indices.argtypes = [C.c_void_p, C.c_int32, u64p]
        L.orc_bitset_indices.restype = C.c_uint64
        L.orc_bitset_bools.argtypes = [C.c_void_p, C.POINTER(C.c_uint8)]
        L.orc_pairs_new.restype = C.c_void_p
        L.orc_pairs_free.argtypes = [C.c_void_p]
        L.orc_pairs_len.argtypes = [C.c_void_p]
        L.orc_pairs_len.restype = C.c_uint64
        L.orc_pairs_build.argtypes = [C.c_void_p]
        L.orc_pairs_build.restype = u64p
        L.orc_pairs_probe.argtypes = [C.c_void_p]
        L.orc_pairs_probe.restype = u32p
        L.orc_pairs_clear.argtypes = [C.c_void_p]
        L.orc_pairs_push.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32]
        L.orc_match_pairs.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.POINTER(_Col), C.POINTER(_Col),
                                      C.c_uint64, u64p, C.c_int32, C.c_void_p]
        L.orc_finish_batch.argtypes = [C.c_int32, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]
        L.orc_finish_batch.restype = C.c_int32
        L.orc_finish_join.argtypes = [C.c_int32, C.c_void_p, C.c_void_p]
        L.orc_finish_join.restype = C.c_int32
        L.orc_make_exponential_int_array.argtypes = [C.c_int32, C.c_int32, i32p]
        L.orc_make_int_array_with_shift.argtypes = [C.c_int32, C.c_int32, C.c_int32, i32p]
        L.orc_gen_permutation_i64.argtypes = [C.c_uint64, C.c_uint64, i64p]
        L.orc_gen_uniform_i64.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, i64p]
        L.mt_table_new.argtypes = [C.c_uint64]
        L.mt_table_new.restype = C.c_void_p
        L.mt_table_free.argtypes = [C.c_void_p]
        L.mt_build.argtypes = [C.c_void_p, i64p, C.c_uint64, u64p, C.c_int32]
        L.mt_build.restype = C.c_double
        L.mt_probe.argtypes = [C.c_void_p, i64p, C.c_uint64, u64p, C.c_int32, u64p, u64p]
        L.mt_probe.restype = C.c_double
        L.mt_max_threads.restype = C.c_int32
        _lib = L
    return _lib


def _p(arr: np.ndarray, ctype):
    return arr.ctypes.data_as(C.POINTER(ctype))


# ------------------------------------------------------------------------------------------------
# hashing
# ------------------------------------------------------------------------------------------------

def keys_with_seed(seed: int = 0) -> np.ndarray:
    """ahash RandomState::with_seed(seed) over the fixed PI seeds (shared.rs:14 uses seed 0)."""
    out = np.zeros(4, dtype=np.uint64)
    lib().orc_keys_with_seed(seed, _p(out, C.c_uint64))
    return out


def hash_u64(keys: np.ndarray, v: int, flags: int = 0) -> int:
    return int(lib().orc_hash_u64(_p(np.ascontiguousarray(keys, dtype=np.uint64), C.c_uint64),
                                  C.c_uint64(v & 0xFFFFFFFFFFFFFFFF), flags))


class _ColSet:
    """Keeps the numpy/arrow buffers alive next to the C descriptors that point into them."""

    def __init__(self, cols: Sequence):
        self.keep = []
        self.n_rows = None
        self.arr = (_Col * max(1, len(cols)))()
        self.n = len(cols)
        for i, col in enumerate(cols):
            offsets = None
            if isinstance(col, pa.ChunkedArray):
                col = col.combine_chunks()
            if isinstance(col, pa.Array) and (pa.types.is_string(col.type) or pa.types.is_binary(col.type)):
                values, offsets, validity, width, n = self._split_utf8(col)
            else:
                values, validity, width, n = self._split(col)
            self.keep += [values, validity, offsets]
            self.arr[i].offsets = offsets.ctypes.data if offsets is not None else None
            if self.n_rows is None:
                self.n_rows = n
            assert self.n_rows == n, "key columns differ in length"
            self.arr[i].values = values.ctypes.data if values.size else None
            self.arr[i].validity = validity.ctypes.data if validity is not None else None
            self.arr[i].width = width
        if self.n_rows is None:
            self.n_rows = 0

    @staticmethod
    def _split_utf8(col):
        """Arrow Utf8: data bytes + n + 1 int32 offsets (rebased to 0); width 0 marks the column as Utf8."""
        n = len(col)
        validity = None
        if col.null_count:
            validity = np.packbits(np.asarray(col.is_valid()).astype(np.uint8), bitorder="little")
        bufs = col.buffers()
        off = np.frombuffer(bufs[1], dtype=np.int32)[col.offset: col.offset + n + 1].copy()
        data = np.frombuffer(bufs[2], dtype=np.uint8)[int(off[0]): int(off[-1])].copy() if bufs[2] is not None and n else np.zeros(0, np.uint8)
        off -= off[0]
        if data.size == 0:
            data = np.zeros(1, np.uint8)
        return data, off, validity, 0, n

    @staticmethod
    def _split(col):
        if isinstance(col, pa.ChunkedArray):
            col = col.combine_chunks()
        if isinstance(col, pa.Array):
            if pa.types.is_int32(col.type) or pa.types.is_uint32(col.type) or pa.types.is_date32(col.type):
                np_t, width = np.uint32, 4
            elif pa.types.is_int64(col.type) or pa.types.is_uint64(col.type):
                np_t, width = np.uint64, 8
            else:
                raise TypeError(f"oracle key type {col.type} unsupported")
            n = len(col)
            validity = None
            if col.null_count:
                validity = np.packbits(np.asarray(col.is_valid()).astype(np.uint8), bitorder="little")
            filled = col.fill_null(0) if col.null_count else col
            values = np.ascontiguousarray(filled.to_numpy(zero_copy_only=False)).view(np_t).copy()
            return values, validity, width, n
        arr = np.ascontiguousarray(col)
        if arr.dtype.itemsize == 4:
            return arr.view(np.uint32), None, 4, arr.shape[0]
        if arr.dtype.itemsize == 8:
            return arr.view(np.uint64), None, 8, arr.shape[0]
        raise TypeError(f"oracle key dtype {arr.dtype} unsupported")


def create_hashes(cols: Sequence, keys: Optional[np.ndarray] = None, flags: int = 0) -> np.ndarray:
    """calculate_hash (shared.rs:11-16): one u64 per row over the key columns."""
    keys = keys_with_seed(0) if keys is None else np.ascontiguousarray(keys, dtype=np.uint64)
    cs = _ColSet(cols)
    out = np.zeros(cs.n_rows, dtype=np.uint64)
    lib().orc_create_hashes(cs.n, cs.arr, cs.n_rows, _p(keys, C.c_uint64), flags, _p(out, C.c_uint64))
    return out


# ------------------------------------------------------------------------------------------------
# chain map / bitset
# ------------------------------------------------------------------------------------------------

class ChainMap:
    """hash -> LIFO list of global row ids (version10 semantics)."""

    def __init__(self):
        self.h = lib().orc_map_new()

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_map_free(self.h)
            self.h = None

    def add(self, hashes) -> None:
        a = np.ascontiguousarray(hashes, dtype=np.uint64)
        lib().orc_map_add(self.h, _p(a, C.c_uint64), a.shape[0])

    insert_all = add

    @property
    def rows(self) -> int:
        return int(lib().orc_map_rows(self.h))

    def get_all(self, hash_value: int) -> List[int]:
        n = int(lib().orc_map_get_all(self.h, C.c_uint64(hash_value), None, 0))
        out = np.zeros(max(n, 1), dtype=np.uint64)
        lib().orc_map_get_all(self.h, C.c_uint64(hash_value), _p(out, C.c_uint64), n)
        return [int(x) for x in out[:n]]


class BitSet:
    def __init__(self, length: int):
        self.length = length
        self.h = lib().orc_bitset_new(length)

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_bitset_free(self.h)
            self.h = None

    def set_ones(self, indices) -> None:
        a = np.ascontiguousarray(indices, dtype=np.uint64)
        if lib().orc_bitset_set_ones(self.h, _p(a, C.c_uint64), a.shape[0]) != 0:
            raise IndexError("Index out of bounds")

    def get_set_count(self) -> int:
        return int(lib().orc_bitset_set_count(self.h))

    def _indices(self, want_set: bool) -> np.ndarray:
        out = np.zeros(max(self.length, 1), dtype=np.uint64)
        n = int(lib().orc_bitset_indices(self.h, int(want_set), _p(out, C.c_uint64)))
        return out[:n].copy()

    def get_set_indices_array(self) -> np.ndarray:
        return self._indices(True)

    def get_unset_indices_array(self) -> np.ndarray:
        return self._indices(False)

    def get_set_indices(self) -> np.ndarray:
        out = np.zeros(max(self.length, 1), dtype=np.uint8)
        lib().orc_bitset_bools(self.h, _p(out, C.c_uint8))
        return out[: self.length].astype(bool)


# ------------------------------------------------------------------------------------------------
# the join
# ------------------------------------------------------------------------------------------------

class _Pairs:
    def __init__(self):
        self.h = lib().orc_pairs_new()

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_pairs_free(self.h)
            self.h = None

    def numpy(self) -> Tuple[np.ndarray, np.ndarray]:
        n = int(lib().orc_pairs_len(self.h))
        if n == 0:
            return np.zeros(0, np.uint64), np.zeros(0, np.uint32)
        b = np.ctypeslib.as_array(lib().orc_pairs_build(self.h), shape=(n,)).copy()
        p = np.ctypeslib.as_array(lib().orc_pairs_probe(self.h), shape=(n,)).copy()
        return b, p

    def assign(self, b: np.ndarray, p: np.ndarray) -> None:
        lib().orc_pairs_clear(self.h)
        for x, y in zip(b.tolist(), p.tolist()):
            lib().orc_pairs_push(self.h, C.c_uint64(x), C.c_uint32(y))


FilterFn = Callable[[np.ndarray, np.ndarray, int], np.ndarray]


class OracleJoin:
    """Build once, probe batch by batch, then finish — the shape of execute_streaming_join
    (parallel_hash_join_executor.rs:86-122) for a single partition."""

    def __init__(self, join_type: str, build_key_cols: Sequence, keys: Optional[np.ndarray] = None,
                 hash_flags: int = 0, build_batch_rows: Optional[int] = None):
        self.join_type = JOIN_TYPES[join_type]
        self.keys = keys_with_seed(0) if keys is None else np.ascontiguousarray(keys, dtype=np.uint64)
        self.flags = hash_flags
        self.build_cols = _ColSet(build_key_cols)
        self.n_build = self.build_cols.n_rows
        self.map = ChainMap()
        hashes = create_hashes(build_key_cols, self.keys, self.flags) if self.n_build else np.zeros(0, np.uint64)
        step = build_batch_rows or max(self.n_build, 1)
        for lo in range(0, self.n_build, step):      # process_input_batch, version10/build_implementation.rs:60-71
            self.map.add(hashes[lo:lo + step])
        self.visited = BitSet(self.n_build)

    def match_pairs(self, probe_key_cols: Sequence) -> Tuple[np.ndarray, np.ndarray]:
        pc = _ColSet(probe_key_cols)
        assert pc.n == self.build_cols.n
        out = _Pairs()
        lib().orc_match_pairs(self.map.h, self.join_type, pc.n, self.build_cols.arr, pc.arr, pc.n_rows,
                              _p(self.keys, C.c_uint64), self.flags, out.h)
        return out.numpy()

    def finish_batch(self, build_idx: np.ndarray, probe_idx: np.ndarray, n_probe: int):
        m, out = _Pairs(), _Pairs()
        m.assign(np.asarray(build_idx, np.uint64), np.asarray(probe_idx, np.uint32))
        rc = lib().orc_finish_batch(self.join_type, m.h, n_probe, self.visited.h, out.h)
        if rc != 0:
            raise RuntimeError("orc_finish_batch failed")
        return out.numpy()

    def probe_batch(self, probe_key_cols: Sequence, n_probe: int, filter_fn: Optional[FilterFn] = None):
        b, p = self.match_pairs(probe_key_cols)
        if filter_fn is not None:                    # apply_join_filter_to_indices, datafusion_private.rs:295-328
            keep = filter_fn(b, p, n_probe)
            b, p = b[keep], p[keep]
        return self.finish_batch(b, p, n_probe)

    def finish(self):
        out = _Pairs()
        if lib().orc_finish_join(self.join_type, self.visited.h, out.h) != 0:
            raise RuntimeError("orc_finish_join failed")
        return out.numpy()


def take_nullable(col, idx: np.ndarray, null_value) -> pa.Array:
    """arrow::compute::take with a NULL-able index array (NULL index -> NULL output)."""
    if isinstance(col, pa.ChunkedArray):
        col = col.combine_chunks()
    mask = idx == null_value
    safe = np.where(mask, 0, idx).astype(np.int64)
    indices = pa.array(safe, type=pa.int64(), mask=mask if mask.any() else None)
    if len(col) == 0:
        return pa.nulls(len(idx), col.type)
    return col.take(indices)


def join_tables(build: pa.Table, probe_batches: Sequence[pa.Table], on: Sequence[Tuple[str, str]],
                join_type: str, filter_fn: Optional[Callable] = None,
                build_batch_rows: Optional[int] = None) -> pa.Table:
    """Whole-join oracle: returns the output table (build columns first, then probe columns;
    semi/anti: one side only — build_join_schema, probe_lookup_implementation.rs:77-94)."""
    jt = JOIN_TYPES[join_type]
    bkeys = [build.column(b) for b, _ in on]
    oj = OracleJoin(join_type, bkeys, build_batch_rows=build_batch_rows)
    left_only = jt in (4, 6)
    right_only = jt in (5, 7)
    names = ([f"l.{n}" for n in build.column_names] if not right_only else []) + \
            ([f"r.{n}" for n in probe_batches[0].column_names] if not left_only else []) if probe_batches else []
    probe_schema = probe_batches[0].schema if probe_batches else None
    pieces = []

    def assemble(bidx, pidx, pb):
        cols = []
        if not right_only:
            cols += [take_nullable(build.column(i), bidx, NULL64) for i in range(build.num_columns)]
        if not left_only:
            if pb is None:
                cols += [pa.nulls(len(bidx), f.type) for f in probe_schema]
            else:
                cols += [take_nullable(pb.column(i), pidx, NULL32) for i in range(pb.num_columns)]
        return pa.Table.from_arrays(cols, names=names)

    for pb in probe_batches:
        pkeys = [pb.column(p) for _, p in on]
        ff = None
        if filter_fn is not None:
            ff = (lambda b, p, n, _pb=pb: filter_fn(build, _pb, b, p))
        bidx, pidx = oj.probe_batch(pkeys, pb.num_rows, ff)
        if not left_only:
            pieces.append(assemble(bidx, pidx, pb))
    bidx, pidx = oj.finish()
    if len(bidx) or not pieces:
        pieces.append(assemble(bidx, pidx, None))
    return pa.concat_tables(pieces) if pieces else None


# ------------------------------------------------------------------------------------------------
# generators + MT baseline
# ------------------------------------------------------------------------------------------------

def make_exponential_int_array(lo: int, hi: int) -> np.ndarray:
    out = np.zeros(max(hi - lo, 0), dtype=np.int32)
    lib().orc_make_exponential_int_array(lo, hi, _p(out, C.c_int32))
    return out


def make_int_array_with_shift(lo: int, hi: int, shift: int) -> np.ndarray:
    out = np.zeros(max(hi - lo, 0), dtype=np.int32)
    lib().orc_make_int_array_with_shift(lo, hi, shift, _p(out, C.c_int32))
    return out


def gen_permutation_i64(n: int, seed: int) -> np.ndarray:
    out = np.zeros(n, dtype=np.int64)
    lib().orc_gen_permutation_i64(n, seed, _p(out, C.c_int64))
    return out


def gen_uniform_i64(n: int, rng: int, seed: int) -> np.ndarray:
    out = np.zeros(n, dtype=np.int64)
    lib().orc_gen_uniform_i64(n, rng, seed, _p(out, C.c_int64))
    return out


class MtJoin:
    """All-host-threads CPU join (cpu_join_mt.c) — the timed CPU baseline."""

    def __init__(self, build_keys: np.ndarray, threads: int = 0):
        self.keys = keys_with_seed(0)
        self.build_keys = np.ascontiguousarray(build_keys, dtype=np.int64)
        self.threads = threads
        self.h = lib().mt_table_new(self.build_keys.shape[0])
        self.build_seconds = float(lib().mt_build(self.h, _p(self.build_keys, C.c_int64), self.build_keys.shape[0],
                                                  _p(self.keys, C.c_uint64), threads))

    def __del__(self):
        if getattr(self, "h", None):
            lib().mt_table_free(self.h)
            self.h = None

    def probe(self, probe_keys: np.ndarray) -> Tuple[float, int, int]:
        pk = np.ascontiguousarray(probe_keys, dtype=np.int64)
        n, cs = C.c_uint64(0), C.c_uint64(0)
        secs = float(lib().mt_probe(self.h, _p(pk, C.c_int64), pk.shape[0], _p(self.keys, C.c_uint64),
                                    self.threads, C.byref(n), C.byref(cs)))
        return secs, int(n.value), int(cs.value)

    @staticmethod
    def max_threads() -> int:
        return int(lib().mt_max_threads())
