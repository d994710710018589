"""Thin Python veneer over the C ABI (include/b200join.h): one `HashJoin` = one b200j_handle.

Everything that computes runs in libb200join.so on the GPU; this module only marshals buffers
(numpy / pyarrow host buffers, torch CUDA tensors or raw device pointers) into `b200j_batch` structs.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np

from . import _native as N

JOIN_TYPES = {"inner": 0, "left": 1, "right": 2, "full": 3, "left_semi": 4, "right_semi": 5, "left_anti": 6, "right_anti": 7}
LAYOUTS = {"auto": 0, "direct": 1, "radix": 2}
MEM_HOST, MEM_DEVICE = 0, 1
OUT_PAIRS, OUT_HOST, OUT_MATCHES_ONLY, OUT_COUNT_ONLY, OUT_BUILD_U32, OUT_GATHER_HINT, OUT_PROBE_ORDER = 1, 2, 4, 8, 16, 32, 64
FLAG_TIMING = 1
NULL_BUILD = np.uint64(0xFFFFFFFFFFFFFFFF)
NULL_PROBE = np.uint32(0xFFFFFFFF)
T_UTF8 = 0


class B200JoinError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"b200join error {code}: {msg}")
        self.code = code


class ColumnRef:
    """One Arrow-layout column: addresses + the Python objects that keep them alive."""

    __slots__ = ("values", "offsets", "validity", "bit_offset", "n_rows", "mem", "type", "keep")

    def __init__(self, values, offsets, validity, bit_offset, n_rows, mem, type_, keep):
        self.values, self.offsets, self.validity = values, offsets, validity
        self.bit_offset, self.n_rows, self.mem, self.type, self.keep = bit_offset, n_rows, mem, type_, keep


def _is_torch(x) -> bool:
    return type(x).__module__.startswith("torch")


def column(x, validity=None) -> ColumnRef:
    """numpy array / torch tensor (CPU or CUDA) / pyarrow Array -> ColumnRef."""
    try:
        import pyarrow as pa
    except ImportError:  # pragma: no cover
        pa = None
    if isinstance(x, ColumnRef):
        return x
    if pa is not None and isinstance(x, pa.ChunkedArray):
        x = x.combine_chunks()
    if pa is not None and isinstance(x, pa.Array):
        t = x.type
        bufs = x.buffers()
        vbuf = bufs[0]
        val_addr = vbuf.address if (vbuf is not None and x.null_count > 0) else None
        if pa.types.is_string(t) or pa.types.is_binary(t):
            off_addr = bufs[1].address + 4 * x.offset
            data_addr = bufs[2].address if bufs[2] is not None else 0
            return ColumnRef(data_addr or None, off_addr, val_addr, x.offset, len(x), MEM_HOST, T_UTF8, (x,))
        if pa.types.is_boolean(t):
            raise TypeError("boolean columns are not supported (bit-packed values)")
        width = t.bit_width // 8
        addr = bufs[1].address + width * x.offset if bufs[1] is not None else None
        return ColumnRef(addr, None, val_addr, x.offset, len(x), MEM_HOST, width, (x,))
    if _is_torch(x):
        assert x.is_contiguous()
        mem = MEM_DEVICE if x.is_cuda else MEM_HOST
        vaddr, keep = None, (x,)
        if validity is not None:
            vaddr = validity.data_ptr() if _is_torch(validity) else validity.ctypes.data
            keep = (x, validity)
        return ColumnRef(x.data_ptr() or None, None, vaddr, 0, x.shape[0], mem, x.element_size(), keep)
    a = np.ascontiguousarray(x)
    vaddr, keep = None, (a,)
    if validity is not None:
        validity = np.ascontiguousarray(validity, dtype=np.uint8)
        vaddr, keep = validity.ctypes.data, (a, validity)
    return ColumnRef(a.ctypes.data if a.size else None, None, vaddr, 0, a.shape[0], MEM_HOST, a.dtype.itemsize, keep)


def make_batch(keys: Sequence, payload: Sequence = ()):
    """-> (Batch struct, keepalive). All columns must live in the same memory space."""
    kcols = [column(k) for k in keys]
    pcols = [column(p) for p in payload]
    cols = kcols + pcols
    n_rows = cols[0].n_rows if cols else 0
    mem = cols[0].mem if cols else MEM_HOST
    for c in cols:
        if c.n_rows != n_rows:
            raise ValueError("columns differ in length")
        if c.mem != mem:
            raise ValueError("columns of one batch must all be host or all be device memory")
    karr = (N.Column * max(1, len(kcols)))()
    parr = (N.Column * max(1, len(pcols)))()
    for arr, src in ((karr, kcols), (parr, pcols)):
        for i, c in enumerate(src):
            arr[i].values, arr[i].offsets, arr[i].validity, arr[i].bit_offset = c.values, c.offsets, c.validity, c.bit_offset
    b = N.Batch()
    b.struct_size = C.sizeof(N.Batch)
    b.mem, b.n_rows, b.n_keys, b.n_payload = mem, n_rows, len(kcols), len(pcols)
    b.keys = karr
    b.payload = parr
    return b, (karr, parr, cols)


class DeviceArray:
    """Zero-copy view of a device buffer owned by a result (torch.as_tensor(x, device='cuda') accepts it)."""

    def __init__(self, ptr: int, n: int, typestr: str, owner):
        self.ptr, self.n, self.owner = ptr, n, owner
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr or 0, False), "version": 3, "strides": None}


class PairResult:
    """Owns one b200j_result until .release()."""

    def __init__(self, join: "HashJoin", res: N.Result):
        self._join, self._res = join, res
        self.n = int(res.n)
        self.mem = int(res.mem)
        self.n_build_null = int(res.n_build_null)
        self.n_probe_null = int(res.n_probe_null)
        self.build_idx_width = int(res.build_idx_width) or 8
        self.probe_order = bool(res.probe_order)      # B200J_OUT_PROBE_ORDER honoured: build_idx[i] belongs to probe row i, no probe_idx array

    @property
    def build_ptr(self) -> int:
        return self._res.build_idx or 0

    @property
    def probe_ptr(self) -> int:
        return self._res.probe_idx or 0

    def numpy(self):
        """(build_idx u64, probe_idx u32) as numpy copies; host results only."""
        if self.mem != MEM_HOST:
            raise ValueError("device result: request OUT_HOST or use .device_arrays()")
        n = self.n
        b = np.full(n, NULL_BUILD, dtype=np.uint64)
        p = np.full(n, NULL_PROBE, dtype=np.uint32)
        if n and self._res.build_idx and self.build_idx_width == 4:
            # B200J_OUT_BUILD_U32: widened here for the callers' convenience (NULL side: UINT32_MAX -> UINT64_MAX)
            b32 = np.ctypeslib.as_array(C.cast(self._res.build_idx, C.POINTER(C.c_uint32)), shape=(n,))
            b = np.where(b32 == NULL_PROBE, NULL_BUILD, b32.astype(np.uint64))
        elif n and self._res.build_idx:
            b = np.ctypeslib.as_array(C.cast(self._res.build_idx, C.POINTER(C.c_uint64)), shape=(n,)).copy()
        if n and self._res.probe_idx:
            p = np.ctypeslib.as_array(C.cast(self._res.probe_idx, C.POINTER(C.c_uint32)), shape=(n,)).copy()
        elif self.probe_order:
            p = np.arange(n, dtype=np.uint32)          # probe-order result: row i of the batch
        return b, p

    def device_arrays(self):
        if self.mem != MEM_DEVICE:
            raise ValueError("host result")
        b = DeviceArray(self._res.build_idx, self.n, "<i8" if self.build_idx_width == 8 else "<i4", self) if self._res.build_idx else None
        p = DeviceArray(self._res.probe_idx, self.n, "<i4", self) if self._res.probe_idx else None
        return b, p

    def release(self):
        if self._res is not None and self._join._h:
            N.lib().b200j_result_release(self._join._h, C.byref(self._res))
        self._res = None

    def __del__(self):
        try:
            self.release()
        except Exception:
            pass


class GatheredResult:
    """Owns one b200j_gathered until .release(). Host results convert to pyarrow arrays (copies)."""

    def __init__(self, join: "HashJoin", g: N.Gathered):
        self._join, self._g = join, g
        self.n_rows, self.n_cols, self.mem = int(g.n_rows), int(g.n_cols), int(g.mem)

    def column_buffers(self, i: int):
        c = self._g.cols[i]
        return {"values": c.values, "offsets": c.offsets, "validity": c.validity, "null_count": int(c.null_count),
                "values_bytes": int(c.values_bytes)}

    def to_arrow(self, types: Sequence) -> list:
        """Host results only: one pyarrow array per output column, `types` = their Arrow types in output order."""
        import pyarrow as pa
        if self.mem != MEM_HOST:
            raise ValueError("device result: request OUT_HOST")
        n, out = self.n_rows, []

        def grab(ptr, nbytes):
            if not ptr or nbytes == 0:
                return pa.py_buffer(b"")
            return pa.py_buffer(C.string_at(ptr, nbytes))

        for i, t in enumerate(types):
            c = self._g.cols[i]
            validity = grab(c.validity, ((n + 31) // 32) * 4) if c.validity else None
            if pa.types.is_string(t) or pa.types.is_binary(t):
                bufs = [validity, grab(c.offsets, (n + 1) * 4), grab(c.values, int(c.values_bytes))]
            else:
                bufs = [validity, grab(c.values, int(c.values_bytes))]
            out.append(pa.Array.from_buffers(t, n, bufs, null_count=int(c.null_count)))
        return out

    def release(self):
        if self._g is not None and self._join._h:
            N.lib().b200j_gathered_release(self._join._h, C.byref(self._g))
        self._g = None

    def __del__(self):
        try:
            self.release()
        except Exception:
            pass


class HashJoin:
    """One join = one b200j_handle: build_push* -> build_finish (per partition) -> probe_batch* -> probe_finish."""

    def __init__(self, join_type="inner", key_widths=(8,), n_partitions=1, device=0, build_payload_types=(),
                 probe_payload_types=(), layout="auto", radix_bits=0, load_factor=0.0, expected_build_rows=0,
                 l2_budget_bytes=0, hash_keys=None, hash_flags=0, stream=None, timing=False):
        L = N.lib()
        self._h = None
        cfg = N.Config()
        cfg.struct_size = C.sizeof(N.Config)
        cfg.join_type = JOIN_TYPES[join_type] if isinstance(join_type, str) else int(join_type)
        cfg.n_partitions, cfg.device = n_partitions, device
        self._kw = (C.c_int32 * max(1, len(key_widths)))(*key_widths)
        self._bp = (C.c_int32 * max(1, len(build_payload_types)))(*build_payload_types)
        self._pp = (C.c_int32 * max(1, len(probe_payload_types)))(*probe_payload_types)
        cfg.n_keys, cfg.key_widths = len(key_widths), self._kw
        cfg.n_build_payload, cfg.build_payload_types = len(build_payload_types), self._bp
        cfg.n_probe_payload, cfg.probe_payload_types = len(probe_payload_types), self._pp
        if hash_keys is not None:
            for i in range(4):
                cfg.hash_keys[i] = int(hash_keys[i])
        cfg.hash_flags = hash_flags
        cfg.layout = LAYOUTS[layout] if isinstance(layout, str) else int(layout)
        cfg.radix_bits, cfg.load_factor = radix_bits, load_factor
        cfg.expected_build_rows, cfg.l2_budget_bytes = expected_build_rows, l2_budget_bytes
        cfg.stream = stream
        cfg.flags = FLAG_TIMING if timing else 0
        h = C.c_void_p()
        rc = L.b200j_create(C.byref(cfg), C.byref(h))
        if rc != 0:
            raise B200JoinError(rc, (L.b200j_last_error(None) or b"").decode())
        self._h = h
        self.join_type = cfg.join_type
        self.n_partitions = n_partitions

    # -- plumbing ---------------------------------------------------------------------------------
    def _check(self, rc: int):
        if rc != 0:
            raise B200JoinError(rc, (N.lib().b200j_last_error(self._h) or b"").decode())

    def close(self):
        if self._h:
            N.lib().b200j_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -- build ------------------------------------------------------------------------------------
    def build_push(self, partition: int, keys: Sequence, payload: Sequence = ()):
        b, keep = make_batch(keys, payload)
        self._check(N.lib().b200j_build_push(self._h, partition, C.byref(b)))

    def build_finish(self, partition: int):
        self._check(N.lib().b200j_build_finish(self._h, partition))

    # -- probe ------------------------------------------------------------------------------------
    def probe_batch(self, partition: int, keys: Sequence, payload: Sequence = (), want: int = OUT_PAIRS | OUT_HOST) -> PairResult:
        b, keep = make_batch(keys, payload)
        r = N.Result()
        r.struct_size = C.sizeof(N.Result)
        self._check(N.lib().b200j_probe_batch(self._h, partition, C.byref(b), want, C.byref(r)))
        return PairResult(self, r)

    def probe_pieces(self, partition: int, pieces: Sequence, want: int = OUT_PAIRS, key_width: int = 0) -> PairResult:
        """One probe batch handed over as several device-resident pieces of its key column: `pieces` = [(device pointer, rows)].
        The batch is their concatenation (b200j_probe_pieces); a pointer may be peer memory mapped with b200j_ipc_open."""
        n = len(pieces)
        ptrs = (C.c_void_p * max(1, n))(*[C.c_void_p(int(p) if p else None) for p, _ in pieces])
        rows = (C.c_int64 * max(1, n))(*[int(m) for _, m in pieces])
        r = N.Result()
        r.struct_size = C.sizeof(N.Result)
        self._check(N.lib().b200j_probe_pieces(self._h, partition, n, ptrs, rows, key_width, want, C.byref(r)))
        return PairResult(self, r)

    # -- coalescing feeder (b200j_probe_push / flush / gather_staged) ----------------------------------
    def probe_push(self, partition: int, keys: Sequence, payload: Sequence = ()):
        b, keep = make_batch(keys, payload)
        self._check(N.lib().b200j_probe_push(self._h, partition, C.byref(b)))

    def staged_rows(self, partition: int) -> int:
        return int(N.lib().b200j_staged_rows(self._h, partition))

    def probe_flush(self, partition: int, want: int = OUT_PAIRS) -> PairResult:
        r = N.Result()
        r.struct_size = C.sizeof(N.Result)
        self._check(N.lib().b200j_probe_flush(self._h, partition, want, C.byref(r)))
        return PairResult(self, r)

    def gather_staged(self, partition: int, pairs: PairResult, want: int = OUT_HOST) -> "GatheredResult":
        g = N.Gathered()
        g.struct_size = C.sizeof(N.Gathered)
        self._check(N.lib().b200j_gather_staged(self._h, partition, C.byref(pairs._res), want, C.byref(g)))
        return GatheredResult(self, g)

    def apply_pairs(self, partition: int, n_probe_rows: int, build_idx, probe_idx, want: int = OUT_PAIRS | OUT_HOST) -> PairResult:
        bi = np.ascontiguousarray(build_idx, dtype=np.uint64)
        pi = np.ascontiguousarray(probe_idx, dtype=np.uint32)
        r = N.Result()
        r.struct_size = C.sizeof(N.Result)
        self._check(N.lib().b200j_apply_pairs(self._h, partition, n_probe_rows, MEM_HOST, bi.shape[0],
                                              bi.ctypes.data if bi.size else None, pi.ctypes.data if pi.size else None,
                                              want, C.byref(r)))
        return PairResult(self, r)

    def apply_pairs_device(self, partition: int, n_probe_rows: int, build_idx_ptr: int, probe_idx_ptr: int, n_pairs: int,
                           want: int = OUT_PAIRS) -> PairResult:
        """b200j_apply_pairs on DEVICE arrays (u64 build indices, u32 probe indices) of caller-filtered candidate pairs."""
        r = N.Result()
        r.struct_size = C.sizeof(N.Result)
        self._check(N.lib().b200j_apply_pairs(self._h, partition, n_probe_rows, MEM_DEVICE, n_pairs, build_idx_ptr or None,
                                              probe_idx_ptr or None, want, C.byref(r)))
        return PairResult(self, r)

    def probe_finish(self, partition: int, want: int = OUT_PAIRS | OUT_HOST) -> PairResult:
        r = N.Result()
        r.struct_size = C.sizeof(N.Result)
        self._check(N.lib().b200j_probe_finish(self._h, partition, want, C.byref(r)))
        return PairResult(self, r)

    # -- gather (a6) ------------------------------------------------------------------------------
    def gather(self, pairs: PairResult, probe_keys: Optional[Sequence] = None, probe_payload: Sequence = (),
               want: int = OUT_HOST) -> "GatheredResult":
        """take_multiple_record_batch (shared.rs:83-92): build key+payload columns at build_idx, then the probe batch's
        key+payload columns at probe_idx. probe_keys=None (the finaliser) yields all-NULL probe columns."""
        g = N.Gathered()
        g.struct_size = C.sizeof(N.Gathered)
        if probe_keys is not None:
            b, keep = make_batch(probe_keys, probe_payload)
            bp = C.byref(b)
        else:
            bp, keep = None, None
        self._check(N.lib().b200j_gather(self._h, C.byref(pairs._res), bp, want, C.byref(g)))
        return GatheredResult(self, g)

    # -- misc -------------------------------------------------------------------------------------
    def hash_keys(self, keys: Sequence) -> np.ndarray:
        b, keep = make_batch(keys)
        if b.mem != MEM_HOST:
            raise ValueError("hash_keys() veneer takes host columns; pass device pointers through the C ABI directly")
        out = np.zeros(b.n_rows, dtype=np.uint64)
        self._check(N.lib().b200j_hash_keys(self._h, C.byref(b), out.ctypes.data if out.size else None))
        return out

    def hash_keys_device(self, keys: Sequence, out_ptr: int):
        b, keep = make_batch(keys)
        self._check(N.lib().b200j_hash_keys(self._h, C.byref(b), out_ptr))

    def stats(self) -> dict:
        s = N.Stats()
        s.struct_size = C.sizeof(N.Stats)
        self._check(N.lib().b200j_get_stats(self._h, C.byref(s)))
        return s.as_dict()

    def reset_stats(self):
        self._check(N.lib().b200j_reset_stats(self._h))

    def synchronize(self):
        self._check(N.lib().b200j_synchronize(self._h))

    def reset(self):
        """Re-arm for another execution of the same join (keeps streams and device allocations)."""
        self._check(N.lib().b200j_reset(self._h))

    def trim(self):
        """Give the handle's cached free device blocks back to the driver (b200j_trim)."""
        self._check(N.lib().b200j_trim(self._h))


class PinnedArray:
    """Page-locked host memory from b200j_host_alloc, viewed as a numpy array (full-speed PCIe copies)."""

    def __init__(self, n: int, dtype):
        self.dtype = np.dtype(dtype)
        self.nbytes = int(n) * self.dtype.itemsize
        p = C.c_void_p()
        rc = N.lib().b200j_host_alloc(max(self.nbytes, 1), C.byref(p))
        if rc != 0:
            raise B200JoinError(rc, (N.lib().b200j_last_error(None) or b"").decode())
        self._p = p
        buf = (C.c_char * max(self.nbytes, 1)).from_address(p.value)
        self.array = np.frombuffer(buf, dtype=self.dtype, count=int(n))

    def free(self):
        if self._p:
            self.array = None
            N.lib().b200j_host_free(self._p)
            self._p = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def device_count() -> int:
    return int(N.lib().b200j_device_count())
