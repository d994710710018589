"""Output ROWS of a join whose sides were handed over as pieces (the multi-GPU exchange): late materialisation across GPUs.

take_multiple_record_batch (src/shared/shared.rs:83-92) needs the build columns at build_idx and the probe columns at
probe_idx. After the exchange those indices count the rows a rank RECEIVED, while the columns stayed on the ranks that own
the rows. The library resolves an index to (piece, row at the piece's source) and reads the value from the source rank's
column through a peer pointer (b200j_resolve_origin, b200j_gather_pieces, b200j_gather_pieces_utf8_*: csrc/gather_remote.cuh).
This module is the ctypes veneer + the bookkeeping of which pointer belongs to which piece; torch only allocates the outputs.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import numpy as np

from . import _native as N
from .join import HashJoin, T_UTF8

NULL_ROW = 0xFFFFFFFF
MAX_PIECES = 64


def _ptrs(ptrs: Sequence[int]):
    return (C.c_void_p * len(ptrs))(*[C.c_void_p(int(p) if p else None) for p in ptrs])


class OriginPieces:
    """The pieces one side of a local join was made of, in the order the local indices count them: for every piece the
    source rank, its row count, where its u32 row ids are (device or peer pointer; 0 = the position inside the piece) and
    what to add to them."""

    def __init__(self):
        self.src: List[int] = []
        self.rows: List[int] = []
        self.rid_ptr: List[int] = []
        self.rid_base: List[int] = []
        self.keep: list = []          # tensors that own local row-id copies

    def add(self, src: int, rows: int, rid_ptr: int, rid_base: int = 0, keep=None):
        self.src.append(int(src))
        self.rows.append(int(rows))
        self.rid_ptr.append(int(rid_ptr or 0))
        self.rid_base.append(int(rid_base))
        if keep is not None:
            self.keep.append(keep)

    def __len__(self):
        return len(self.src)

    @property
    def total(self) -> int:
        return int(sum(self.rows))


class SharedColumn:
    """One column of one join side as every rank can read it: per rank the pointers of its buffers (cudaIpc-mapped for the
    peers). type_ = byte width (1/2/4/8/16) or T_UTF8 (values = int32 offsets, data = the bytes)."""

    def __init__(self, type_: int, values: Sequence[int], validity: Optional[Sequence[int]] = None, data: Optional[Sequence[int]] = None,
                 keep=None):
        self.type = int(type_)
        self.values = [int(p or 0) for p in values]
        self.validity = [int(p or 0) for p in validity] if validity is not None else [0] * len(self.values)
        self.data = [int(p or 0) for p in data] if data is not None else [0] * len(self.values)
        self.keep = keep


def resolve_origin(join: HashJoin, idx_ptr: int, n: int, idx_width: int, origin: OriginPieces, out_piece_ptr: int, out_row_ptr: int):
    """b200j_resolve_origin: local indices -> (piece, row at the piece's source); NULL indices -> row 0xffffffff."""
    m = len(origin)
    if not 1 <= m <= MAX_PIECES:
        raise ValueError(f"1..{MAX_PIECES} pieces per call, got {m}")
    rows = (C.c_int64 * m)(*origin.rows)
    base = (C.c_uint32 * m)(*origin.rid_base)
    join._check(N.lib().b200j_resolve_origin(join._h, n, idx_ptr or None, idx_width, m, rows, _ptrs(origin.rid_ptr), base,
                                             out_piece_ptr or None, out_row_ptr or None))


def gather_pieces(join: HashJoin, piece_ptr: int, row_ptr: int, n: int, values: Sequence[int], validity: Optional[Sequence[int]], width: int,
                  out_ptr: int, out_valid_ptr: int = 0):
    m = len(values)
    join._check(N.lib().b200j_gather_pieces(join._h, n, piece_ptr or None, row_ptr or None, m, _ptrs(values),
                                            _ptrs(validity) if validity is not None else None, width, out_ptr or None, out_valid_ptr or None))


class GatheredColumn:
    """One gathered output column on the device: fixed width -> values (torch tensor of `n` elements; 16-byte values as
    [n, 2] int64) + validity words; Utf8 -> offsets (n + 1 int32), data (uint8), validity words (bit i%32 of word i/32)."""

    def __init__(self, type_: int, n: int, values, validity, offsets=None):
        self.type, self.n, self.values, self.validity, self.offsets = type_, n, values, validity, offsets

    def validity_numpy(self) -> np.ndarray:
        w = self.validity.cpu().numpy().view(np.uint32)
        bits = np.unpackbits(w.view(np.uint8), bitorder="little")[: self.n]
        return bits.astype(bool)

    def to_arrow(self, typ):
        import pyarrow as pa
        valid = self.validity_numpy()
        vbuf = pa.py_buffer(np.packbits(valid, bitorder="little").tobytes()) if self.n else None
        if self.type == T_UTF8:
            off = self.offsets.cpu().numpy().astype(np.int32)
            data = self.values.cpu().numpy().tobytes()
            return pa.Array.from_buffers(typ, self.n, [vbuf, pa.py_buffer(off.tobytes()), pa.py_buffer(data)])
        raw = self.values.cpu().numpy().tobytes()
        return pa.Array.from_buffers(typ, self.n, [vbuf, pa.py_buffer(raw)])


_TORCH_OF_WIDTH = {1: "uint8", 2: "int16", 4: "int32", 8: "int64"}


class RowGatherer:
    """Resolves the indices of ONE side once and gathers any number of that side's columns."""

    def __init__(self, join: HashJoin, torch, idx, n: int, idx_width: int, origin: OriginPieces, stream=None):
        self.join, self.torch, self.n, self.origin = join, torch, int(n), origin
        dev = torch.device("cuda", torch.cuda.current_device())
        self.dev = dev
        self.piece = torch.empty(max(self.n, 1), dtype=torch.int32, device=dev)
        self.row = torch.empty(max(self.n, 1), dtype=torch.int32, device=dev)
        self._keep = idx
        if self.n:
            if len(origin) == 0:
                self.row.fill_(-1)                                # no piece at all: every row is NULL
                self.piece.zero_()
            else:
                torch.cuda.current_stream().synchronize()         # the outputs were just allocated on torch's stream; the library uses its own
                ptr = idx if isinstance(idx, int) else idx.data_ptr() if hasattr(idx, "data_ptr") else int(idx.__cuda_array_interface__["data"][0])
                resolve_origin(join, ptr, self.n, idx_width, origin, self.piece.data_ptr(), self.row.data_ptr())

    @classmethod
    def all_null(cls, join: HashJoin, torch, n: int):
        """The side a finaliser / an unmatched row does not have: every gathered value is NULL."""
        return cls(join, torch, None, n, 4, OriginPieces())

    def _per_piece(self, per_rank: Sequence[int]) -> List[int]:
        if len(self.origin) == 0:
            return [0]
        return [per_rank[s] for s in self.origin.src]

    def gather(self, col: SharedColumn) -> GatheredColumn:
        torch, n = self.torch, self.n
        words = torch.zeros((n + 31) // 32 + 1, dtype=torch.int32, device=self.dev)
        values = self._per_piece(col.values)
        validity = self._per_piece(col.validity)
        if col.type != T_UTF8:
            w = col.type
            out = (torch.empty((max(n, 1), 2), dtype=torch.int64, device=self.dev) if w == 16
                   else torch.empty(max(n, 1), dtype=getattr(torch, _TORCH_OF_WIDTH[w]), device=self.dev))
            torch.cuda.current_stream().synchronize()
            gather_pieces(self.join, self.piece.data_ptr(), self.row.data_ptr(), n, values, validity, w, out.data_ptr(), words.data_ptr())
            self.join.synchronize()
            return GatheredColumn(w, n, out[:n], words)
        data = self._per_piece(col.data)
        m = len(values)
        off = torch.zeros(n + 1, dtype=torch.int32, device=self.dev)
        total = C.c_int64(0)
        torch.cuda.current_stream().synchronize()
        L = N.lib()
        self.join._check(L.b200j_gather_pieces_utf8_offsets(self.join._h, n, self.piece.data_ptr(), self.row.data_ptr(), m, _ptrs(values),
                                                            _ptrs(validity), off.data_ptr(), words.data_ptr(), C.byref(total)))
        out = torch.empty(max(int(total.value), 1), dtype=torch.uint8, device=self.dev)
        torch.cuda.current_stream().synchronize()
        self.join._check(L.b200j_gather_pieces_utf8_data(self.join._h, n, self.piece.data_ptr(), self.row.data_ptr(), m, _ptrs(values),
                                                         _ptrs(data), off.data_ptr(), out.data_ptr()))
        self.join.synchronize()
        return GatheredColumn(T_UTF8, n, out[: int(total.value)], words, offsets=off)


def utf8_rows_equal(join: HashJoin, torch, a: GatheredColumn, b: GatheredColumn):
    """b200j_utf8_rows_equal on two gathered Utf8 columns of the same length -> uint8 tensor (1 = same bytes, both valid)."""
    n = a.n
    out = torch.zeros(max(n, 1), dtype=torch.uint8, device=a.offsets.device)
    torch.cuda.current_stream().synchronize()
    join._check(N.lib().b200j_utf8_rows_equal(join._h, n, a.offsets.data_ptr(), a.values.data_ptr() if a.values.numel() else None,
                                              a.validity.data_ptr(), b.offsets.data_ptr(), b.values.data_ptr() if b.values.numel() else None,
                                              b.validity.data_ptr(), out.data_ptr()))
    join.synchronize()
    return out[:n]


def valid_mask(torch, col: GatheredColumn):
    """bool tensor [n] from the validity words."""
    n = col.n
    i = torch.arange(n, device=col.validity.device)
    return ((col.validity[(i >> 5)] >> (i & 31)) & 1).to(torch.bool)


def arrow_to_device(torch, arr, dev=None):
    """pyarrow Array (fixed width or Utf8, any offset, NULLs) -> dict of device tensors laid out for SharedColumn:
    {"type", "values" (fixed: raw values; Utf8: int32 offsets rebased to 0), "data" (Utf8), "validity" (uint8 bitmap from
    row 0, or None), "n"}."""
    import pyarrow as pa
    if isinstance(arr, pa.ChunkedArray):
        arr = arr.combine_chunks()
    dev = dev or torch.device("cuda", torch.cuda.current_device())
    n = len(arr)
    out = {"n": n, "validity": None, "data": None}
    if arr.null_count:
        valid = np.asarray(arr.is_valid())
        out["validity"] = torch.from_numpy(np.packbits(valid, bitorder="little").copy()).to(dev)
    if pa.types.is_string(arr.type) or pa.types.is_binary(arr.type):
        bufs = arr.buffers()
        off = np.frombuffer(bufs[1], dtype=np.int32, count=n + 1, offset=arr.offset * 4).astype(np.int64)
        data = np.frombuffer(bufs[2], dtype=np.uint8)[int(off[0]): int(off[-1])] if bufs[2] is not None and n else np.zeros(0, np.uint8)
        out["type"] = T_UTF8
        out["values"] = torch.from_numpy((off - off[0]).astype(np.int32)).to(dev)
        out["data"] = torch.from_numpy(data.copy() if data.size else np.zeros(1, np.uint8)).to(dev)
        return out
    width = arr.type.bit_width // 8
    raw = np.frombuffer(arr.buffers()[1], dtype=np.uint8, count=n * width, offset=arr.offset * width) if n else np.zeros(0, np.uint8)
    out["type"] = width
    out["values"] = torch.from_numpy(raw.copy() if raw.size else np.zeros(width, np.uint8)).to(dev)
    return out


def equal_key_rows(join: HashJoin, torch, gb: RowGatherer, gp: RowGatherer, build_keys: Sequence[SharedColumn],
                   probe_keys: Sequence[SharedColumn]):
    """equal_rows_arr (src/shared/datafusion_private.rs:40-80, null_equals_null = false) for candidate pairs of a join that was
    keyed on ROW HASHES (composite / Utf8 keys through the exchange): gathers every key column of both sides from the ranks
    that own the rows and compares them. Returns a bool tensor [n]: the pair's keys are equal and none of them is NULL."""
    n = gb.n
    keep = torch.ones(n, dtype=torch.bool, device=gb.dev)
    for bc, pc in zip(build_keys, probe_keys):
        if bc.type != pc.type:
            raise ValueError("key column types differ between the sides")
        a, b = gb.gather(bc), gp.gather(pc)
        if bc.type == T_UTF8:
            keep &= utf8_rows_equal(join, torch, a, b).to(torch.bool)
        else:
            eq = (a.values == b.values)
            if eq.dim() == 2:
                eq = eq.all(dim=1)
            keep &= eq & valid_mask(torch, a) & valid_mask(torch, b)
    return keep
