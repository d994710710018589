"""Multi-GPU hash join: one process per GPU, both join sides hash-partitioned by destination rank and exchanged
over NVLink, then every GPU joins its own partition locally (the reference is single-process, so this
has no counterpart there).

torch.distributed is the plumbing only (count exchange, barriers, the NCCL all-to-all of mode "nccl"). The
partitioning and, in mode "p2p", the exchange itself are libb200join kernels: b200j_shuffle_scatter writes every
(key, row id) straight into the destination GPU's receive buffer through cudaIpc-mapped peer pointers, so the
scatter kernel IS the all-to-all (no send buffer, no second pass over the data).

Host-side planning (`ExchangePlan`) is pure Python over the count matrix, so it is unit-tested on CPU with gloo.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import numpy as np

from . import _native as N
from .join import HashJoin, B200JoinError, MEM_DEVICE, OUT_PAIRS, make_batch

IPC_BYTES = 64


class ExchangePlan:
    """What every rank derives from the all-gathered count matrix counts[src][dst] (rows of src that hash to dst)."""

    def __init__(self, counts: np.ndarray, rank: int, src_order: Optional[Sequence[int]] = None):
        counts = np.asarray(counts, dtype=np.int64)
        world = counts.shape[0]
        assert counts.shape == (world, world)
        self.world, self.rank, self.counts = world, rank, counts
        # the order in which the sources' rows sit in MY received batch: rank order for the push modes; rotated to start behind
        # me in mode "pull" (all ranks reading source 0 first would queue up on one GPU's NVLink egress)
        self.src_order = np.arange(world) if src_order is None else np.asarray(src_order, dtype=np.int64)
        self.recv_rows = counts.sum(axis=0)                      # rows every rank receives
        self.max_recv = int(self.recv_rows.max())
        # where MY rows start inside destination d's receive buffer: after the rows of lower ranks
        self.dest_offset = np.array([int(counts[:rank, d].sum()) for d in range(world)], dtype=np.int64)
        # where the rows of the k-th source (src_order[k]) start inside MY received batch
        self.src_offset = np.concatenate([[0], np.cumsum(counts[self.src_order, rank])]).astype(np.int64)
        self.send_split = [int(x) for x in counts[rank, :]]
        self.recv_split = [int(x) for x in counts[:, rank]]
        self.send_offset = np.concatenate([[0], np.cumsum(counts[rank, :])]).astype(np.int64)

    def origin_of(self, recv_index: np.ndarray):
        """(source rank, position inside that source's segment) of rows of my receive buffer."""
        k = np.searchsorted(self.src_offset[1:], recv_index, side="right")
        return self.src_order[k], recv_index - self.src_offset[k]


class _PeerBuffer:
    """A device buffer every rank of the node can write: cudaMalloc + cudaIpc handle, opened by all peers."""

    def __init__(self, join: HashJoin, dist, nbytes: int):
        L = N.lib()
        self.join, self.nbytes = join, nbytes
        p = C.c_void_p()
        hbuf = C.create_string_buffer(IPC_BYTES)
        join._check(L.b200j_ipc_alloc(join._h, nbytes, C.byref(p), hbuf))
        self.local = p.value
        world, rank = dist.get_world_size(), dist.get_rank()
        handles: List[Optional[bytes]] = [None] * world
        dist.all_gather_object(handles, hbuf.raw)
        self.ptrs: List[int] = []
        self._opened: List[int] = []
        for r in range(world):
            if r == rank:
                self.ptrs.append(self.local)
                continue
            q = C.c_void_p()
            join._check(L.b200j_ipc_open(join._h, handles[r], C.byref(q)))
            self.ptrs.append(q.value)
            self._opened.append(q.value)

    def close(self, dist):
        L = N.lib()
        dist.barrier()
        for q in self._opened:
            L.b200j_ipc_close(self.join._h, q)
        self._opened = []
        dist.barrier()
        if self.local:
            L.b200j_ipc_free(self.join._h, self.local)
            self.local = None


def _dev_column(ptr: int, n_rows: int, width: int, keep=None):
    from .join import ColumnRef
    return ColumnRef(ptr or None, None, None, 0, n_rows, MEM_DEVICE, width, keep)


def shuffle_set_hot(join: HashJoin, hot_keys) -> None:
    """b200j_shuffle_set_hot: key values whose rows stay on the rank that holds them (empty / None = off)."""
    hk = np.ascontiguousarray(np.asarray(hot_keys if hot_keys is not None else [], dtype=np.uint64))
    ptr = hk.ctypes.data_as(C.POINTER(C.c_uint64)) if hk.size else None
    join._check(N.lib().b200j_shuffle_set_hot(join._h, ptr, int(hk.size)))


def shuffle_count(join: HashJoin, keys, n_dest: int, hot: bool = False) -> np.ndarray:
    """b200j_shuffle_count: rows of `keys` (device column) per destination rank; [n_dest] = NULL-key rows,
    [n_dest + 1] = rows of hot keys (only with a hot set)."""
    b, keep = make_batch([keys])
    counts = (C.c_int64 * (n_dest + 2))()
    join._check(N.lib().b200j_shuffle_count(join._h, C.byref(b), n_dest, counts))
    return np.array(list(counts), dtype=np.int64)


STATIC_MAX_ROWS = 1 << 27        # B200J_STATIC_MAX_ROWS


def shuffle_count_static(join: HashJoin, keys_ptr: int, n_rows: int, n_dest: int, counts_ptr: int, validity_ptr: int = 0,
                         bit_offset: int = 0) -> None:
    """b200j_shuffle_count_static: the counts of shuffle_count, written to the caller's DEVICE buffer `counts_ptr`
    ([n_dest + 1] int64) by work enqueued on the handle's stream. Allocates nothing (the kernel's scratch is a static device
    array of the library); at most STATIC_MAX_ROWS rows per call; calls that overlap on the GPU race on that array."""
    join._check(N.lib().b200j_shuffle_count_static(join._h, keys_ptr or None, validity_ptr or None, bit_offset, n_rows, n_dest,
                                                   counts_ptr or None))


def shuffle_scatter(join: HashJoin, n_dest: int, dest_keys: Sequence[int], dest_rowid: Sequence[int], dest_offset: Sequence[int],
                    rowid_base: int = 0):
    """b200j_shuffle_scatter: write every (key, row id) of the counted batch to its destination buffer."""
    m = len(dest_keys)                      # n_dest + 1, or n_dest + 2 with a hot set
    dk = (C.c_void_p * m)(*[C.c_void_p(int(p) if p else None) for p in dest_keys])
    dr = (C.c_void_p * m)(*[C.c_void_p(int(p) if p else None) for p in dest_rowid])
    do = (C.c_int64 * m)(*[int(x) for x in dest_offset])
    join._check(N.lib().b200j_shuffle_scatter(join._h, n_dest, dk, dr, do, rowid_base))


def shuffle_local_scatter(join: HashJoin, keys, n_dest: int, send_keys_ptr: int, send_rid_ptr: int, region_rows: int, rowid_base: int = 0,
                          key_out_width: int = 0) -> np.ndarray:
    """b200j_shuffle_local_scatter: one pass over `keys` (device column); every (key, row id) lands in the region of its
    destination inside the local send buffers. Returns the rows per region: [ranks][NULL keys]([hot keys])."""
    b, keep = make_batch([keys])
    counts = (C.c_int64 * (n_dest + 2))()
    join._check(N.lib().b200j_shuffle_local_scatter(join._h, C.byref(b), n_dest, send_keys_ptr, send_rid_ptr, region_rows, rowid_base,
                                                    key_out_width, counts))
    return np.array(list(counts), dtype=np.int64)


def copy_async(join: HashJoin, dst: int, src: int, nbytes: int, lane: int = 0):
    join._check(N.lib().b200j_copy_async(join._h, dst, src, nbytes, lane))


def copy_wait(join: HashJoin):
    join._check(N.lib().b200j_copy_wait(join._h))


class _Side:
    """Receive (and, for NCCL, send) buffers of one join side on this rank."""

    def __init__(self):
        self.cap = 0
        self.keys = None        # _PeerBuffer (p2p) or torch tensor (nccl)
        self.rid = None
        self.send_keys = None   # nccl only
        self.send_rid = None
        self.plan: Optional[ExchangePlan] = None
        self.null_keys = None
        self.null_rid = None
        self.hot_keys = None    # rows of hot keys, kept on this rank (heavy-hitter routing)
        self.hot_rid = None
        self.n_hot = 0


class DistributedJoin:
    """HashJoin over `world` GPUs. Same call shape as HashJoin for one partition per rank; inputs are device tensors."""

    def __init__(self, join_type="inner", key_width=8, stream=None, mode="p2p", group=None, **join_kwargs):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.group = group
        self.world, self.rank = dist.get_world_size(), dist.get_rank()
        self.mode, self.kw = mode, key_width
        self.track_origin = join_kwargs.pop("track_origin", False)   # keep (source rank, row id) of every build row for build_rows()
        self.narrow_keys = join_kwargs.pop("narrow_keys", True)      # mode "pull": 8-byte probe keys cross NVLink as 32-bit values while they fit
        self.stream = stream if stream is not None else torch.cuda.current_stream()
        join_kwargs.setdefault("device", torch.cuda.current_device())     # one process per GPU: the handle lives on this rank's device
        self.join = HashJoin(join_type, key_widths=(key_width,), stream=self.stream.cuda_stream, **join_kwargs)
        # a second handle on its own stream does the partition + exchange kernels of probe chunk k+1 while the first one
        # joins chunk k (probe_chunks): two receive-buffer sets for the probe side
        self.xstream = torch.cuda.Stream(device=join_kwargs["device"])
        xkw = {k: v for k, v in join_kwargs.items() if k in ("device", "hash_keys", "hash_flags")}
        self.xjoin = HashJoin(join_type, key_widths=(key_width,), stream=self.xstream.cuda_stream, **xkw)
        self.build = _Side()
        self.probe_sides = [_Side(), _Side()]
        self.probe = self.probe_sides[0]
        self.ms_exchange = {"build": 0.0, "probe": 0.0}
        self.join_type = join_type
        self.hot = None                 # numpy uint64 hot key values, or None
        self.hot_join: Optional[HashJoin] = None
        self.hot_build_keys = self.hot_owner = self.hot_bidx = None
        self.build_rows_local = 0
        self.build_rows_sent = 0        # rows of this rank's build batches so far: the row-id base of the next batch
        self.build_null_rid = []        # row ids of this rank's NULL-key build rows (never sent), per build batch
        self.build_origin = []          # OriginPieces per build batch received (row-id copies: where every local build row came from)
        self._key_dtype = torch.int64 if key_width == 8 else torch.int32

    # ---- buffers ----------------------------------------------------------------------------------------------
    def _ensure(self, side: _Side, rows: int):
        """Collective: every rank calls with the same `rows` (the max over ranks of what anyone receives)."""
        if rows <= side.cap:
            return
        torch = self.torch
        cap = int(rows * 1.1) + 4096
        if self.mode in ("p2p", "ce"):
            for b in (side.keys, side.rid):
                if b is not None:
                    b.close(self.dist)
            # allocated through the exchange handle: the join handle may be busy with a local probe (probe_chunks)
            side.keys = _PeerBuffer(self.xjoin, self.dist, cap * self.kw + 64)
            side.rid = _PeerBuffer(self.xjoin, self.dist, cap * 4 + 64)
        else:
            dev = torch.device("cuda", torch.cuda.current_device())
            side.keys = torch.empty(cap, dtype=self._key_dtype, device=dev)
            side.rid = torch.empty(cap, dtype=torch.int32, device=dev)
        side.cap = cap

    def _exchange(self, side: _Side, keys, rowid_base: int = 0, overlapped: bool = False) -> int:
        """Partition `keys` by destination rank and deliver them; returns the rows this rank received.
        overlapped: run on the exchange handle / stream (called from the exchange thread of probe_chunks)."""
        torch, dist = self.torch, self.dist
        W = self.world
        n = int(keys.shape[0])
        xj = self.xjoin if overlapped else self.join
        xs = self.xstream if overlapped else self.stream
        hot = self.hot is not None and side is not self.build        # the build side is never routed: its rows go to their owners
        if self.hot is not None:
            shuffle_set_hot(xj, self.hot if hot else None)
        if self.mode == "ce":
            n_recv = self._exchange_ce(side, keys, rowid_base, xj, xs, hot)
            side.pieces = [(side.recv_keys_ptr, n_recv)]
            side.rid_pieces = [(side.recv_rid_ptr, n_recv)]
            return n_recv
        if self.mode == "pull":
            return self._exchange_pull(side, keys, rowid_base, xj, xs, hot)
        counts = shuffle_count(xj, keys, W, hot)
        with torch.cuda.stream(xs):
            mine = torch.from_numpy(counts[:W].copy()).to(keys.device)
            allc = torch.empty(W * W, dtype=torch.int64, device=keys.device)
            dist.all_gather_into_tensor(allc, mine, group=self.group)
            matrix = allc.cpu().numpy().reshape(W, W)
        plan = ExchangePlan(matrix, self.rank)
        side.plan = plan
        self._ensure(side, plan.max_recv)
        n_null = int(counts[W])
        nk = nr = 0
        if n_null:
            side.null_keys = torch.empty(n_null, dtype=self._key_dtype, device=keys.device)
            side.null_rid = torch.empty(n_null, dtype=torch.int32, device=keys.device)
            nk, nr = side.null_keys.data_ptr(), side.null_rid.data_ptr()
        extra_k, extra_r, extra_o = [nk], [nr], [0]
        side.n_hot = 0
        if hot:
            side.n_hot = int(counts[W + 1])
            side.hot_keys = torch.empty(max(side.n_hot, 1), dtype=self._key_dtype, device=keys.device)
            side.hot_rid = torch.empty(max(side.n_hot, 1), dtype=torch.int32, device=keys.device)
            extra_k, extra_r, extra_o = [nk, side.hot_keys.data_ptr()], [nr, side.hot_rid.data_ptr()], [0, 0]
        if self.mode == "p2p":
            # the count all-gather above doubles as the "everyone is done with the previous contents" barrier
            shuffle_scatter(xj, W, list(side.keys.ptrs) + extra_k, list(side.rid.ptrs) + extra_r,
                            list(plan.dest_offset) + extra_o, rowid_base)
            xj.synchronize()
            dist.barrier(group=self.group)          # every peer's stores have landed in my receive buffer
            recv_keys_ptr, recv_rid_ptr = side.keys.local, side.rid.local
        else:
            if side.send_keys is None or side.send_keys.shape[0] < n:
                side.send_keys = torch.empty(int(n * 1.1) + 4096, dtype=self._key_dtype, device=keys.device)
                side.send_rid = torch.empty(int(n * 1.1) + 4096, dtype=torch.int32, device=keys.device)
            sk, sr = side.send_keys.data_ptr(), side.send_rid.data_ptr()
            shuffle_scatter(xj, W, [sk] * W + extra_k, [sr] * W + extra_r, list(plan.send_offset[:W]) + extra_o, rowid_base)
            n_recv = int(plan.recv_rows[self.rank])
            n_send = int(plan.send_offset[W])
            with torch.cuda.stream(xs):
                dist.all_to_all_single(side.keys[:n_recv], side.send_keys[:n_send], plan.recv_split, plan.send_split, group=self.group)
                dist.all_to_all_single(side.rid[:n_recv], side.send_rid[:n_send], plan.recv_split, plan.send_split, group=self.group)
            xs.synchronize()
            recv_keys_ptr, recv_rid_ptr = side.keys.data_ptr(), side.rid.data_ptr()
        side.recv_keys_ptr, side.recv_rid_ptr = recv_keys_ptr, recv_rid_ptr
        n_recv = int(plan.recv_rows[self.rank])
        side.pieces = [(recv_keys_ptr, n_recv)]
        side.rid_pieces = [(recv_rid_ptr, n_recv)]
        return n_recv

    def _exchange_ce(self, side: _Side, keys, rowid_base: int, xj: HashJoin, xs, hot: bool) -> int:
        """Mode "ce": the exchange is carried by the copy engines. ONE kernel scatters the batch by destination into local send
        regions (single read of the keys, no histogram pass: b200j_shuffle_local_scatter), the ranks all-gather the counts,
        and every region then crosses NVLink as one cudaMemcpyAsync into the peer's receive buffer at its exact (dense)
        offset -- no SM is busy with the transfer, so it overlaps the local join of the previous chunk for real (the fused
        peer-store scatter of mode "p2p" competes with the join for SMs and L2: 33 -> 40 ms/step at 2 GPUs when overlapped)."""
        torch, dist = self.torch, self.dist
        W = self.world
        n = int(keys.shape[0])
        n_regions = W + 2
        grow = 1.0
        while True:
            need = int(n / W * 1.15 * grow) + 65536
            if side.send_keys is None or getattr(side, "send_cap", 0) < need:
                side.send_keys = torch.empty(n_regions * need, dtype=self._key_dtype, device=keys.device)
                side.send_rid = torch.empty(n_regions * need, dtype=torch.int32, device=keys.device)
                side.send_cap = need
            cap = side.send_cap
            try:
                counts, ovf = shuffle_local_scatter(xj, keys, W, side.send_keys.data_ptr(), side.send_rid.data_ptr(), cap, rowid_base), 0
            except B200JoinError as e:
                if "region too small" not in str(e):
                    raise
                counts, ovf = np.zeros(W + 2, dtype=np.int64), 1
            # every rank has to take the same branch: the overflow flag travels with the counts
            with torch.cuda.stream(xs):
                mine = torch.from_numpy(np.concatenate([counts[:W], [ovf]]).astype(np.int64)).to(keys.device)
                allc = torch.empty(W * (W + 1), dtype=torch.int64, device=keys.device)
                dist.all_gather_into_tensor(allc, mine, group=self.group)
                full = allc.cpu().numpy().reshape(W, W + 1)
            if int(full[:, W].sum()) == 0:
                break
            if ovf:                                 # a skewed destination: twice the region size on my retry
                grow *= 2.0
        matrix = full[:, :W]
        plan = ExchangePlan(matrix, self.rank)
        side.plan = plan
        self._ensure(side, plan.max_recv)
        kw = self.kw
        sk, sr = side.send_keys.data_ptr(), side.send_rid.data_ptr()
        for i in range(W):
            d = (self.rank + i) % W             # start with my own region: the ranks do not all hit rank 0 first
            m = int(counts[d])
            if m == 0:
                continue
            off = int(plan.dest_offset[d])
            copy_async(xj, side.keys.ptrs[d] + off * kw, sk + d * cap * kw, m * kw, lane=(2 * i) % 4)
            copy_async(xj, side.rid.ptrs[d] + off * 4, sr + d * cap * 4, m * 4, lane=(2 * i + 1) % 4)
        n_null = int(counts[W])
        side.null_keys = side.send_keys[W * cap: W * cap + n_null] if n_null else None
        side.null_rid = side.send_rid[W * cap: W * cap + n_null] if n_null else None
        side.n_hot = 0
        if hot:
            side.n_hot = int(counts[W + 1])
            side.hot_keys = side.send_keys[(W + 1) * cap: (W + 1) * cap + max(side.n_hot, 1)]
            side.hot_rid = side.send_rid[(W + 1) * cap: (W + 1) * cap + max(side.n_hot, 1)]
        copy_wait(xj)
        dist.barrier(group=self.group)          # every peer's copies have landed in my receive buffer
        side.recv_keys_ptr, side.recv_rid_ptr = side.keys.local, side.rid.local
        return int(plan.recv_rows[self.rank])

    def _exchange_pull(self, side: _Side, keys, rowid_base: int, xj: HashJoin, xs, hot: bool) -> int:
        """Mode "pull": nothing is pushed. ONE kernel scatters the batch by destination into LOCAL send regions that every peer
        has mapped (cudaIpc); the ranks all-gather the counts; the destination's local join then takes the W regions meant for
        it as the PIECES of one batch (b200j_probe_pieces / build_push per piece) and its partition pass pulls the keys over
        NVLink while it partitions them: the transfer is fused into a pass the join makes anyway, only the 8-byte keys cross
        the link (the source row ids stay where they are until somebody asks for an origin), and no receive buffer is written
        and read back. Loads can be kept in flight by the thousands per SM, peer STORES cannot (the push scatter of mode "p2p"
        reaches 480-550 GB/s of the 770 GB/s a peer copy gets)."""
        torch, dist = self.torch, self.dist
        W, kw = self.world, self.kw
        n = int(keys.shape[0])
        n_regions = W + 2
        grow = 1.0
        first = True
        while True:
            need = (int(n / W * 1.15 * grow) + 65536 + 63) // 64 * 64
            if first or grow > 1.0:
                # ONE collective: region size = the largest any rank needs, and nobody is still pulling from the regions about to be
                # overwritten (every rank gets here only after its previous local join has returned)
                with torch.cuda.stream(xs):
                    t = torch.tensor([need], dtype=torch.int64, device=keys.device)
                    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
                    need = int(t.item())
                first = False
            if getattr(side, "send_cap", 0) < need:       # collective: every rank sees the same `need`
                for b in (getattr(side, "send_pk", None), getattr(side, "send_pr", None)):
                    if b is not None:
                        b.close(dist)
                side.send_pk = _PeerBuffer(xj, dist, n_regions * need * kw + 64)
                side.send_pr = _PeerBuffer(xj, dist, n_regions * need * 4 + 64)
                side.send_cap = need
            cap = side.send_cap
            # probe-side keys cross NVLink as 32-bit values while every rank's keys fit (8-byte keys only; the build side is
            # appended to the store as it is, so it stays wide)
            okw = 4 if (kw == 8 and side is not self.build and self.narrow_keys and getattr(side, "narrow", True)) else kw
            ovf = wide = 0
            try:
                counts = shuffle_local_scatter(xj, keys, W, side.send_pk.local, side.send_pr.local, cap, rowid_base, key_out_width=okw)
            except B200JoinError as e:
                counts = np.zeros(W + 2, dtype=np.int64)
                if "region too small" in str(e):
                    ovf = 1
                elif "do not fit 32 bits" in str(e):
                    wide = 1
                else:
                    raise
            with torch.cuda.stream(xs):
                mine = torch.from_numpy(np.concatenate([counts[:W], [ovf, wide]]).astype(np.int64)).to(keys.device)
                allc = torch.empty(W * (W + 2), dtype=torch.int64, device=keys.device)
                dist.all_gather_into_tensor(allc, mine, group=self.group)     # ... and every peer's scatter has finished
                full = allc.cpu().numpy().reshape(W, W + 2)
            if int(full[:, W + 1].sum()):
                side.narrow = False                      # somebody's keys need 64 bits: every rank scatters again, wide, from now on
                continue
            if int(full[:, W].sum()) == 0:
                break
            grow *= 2.0                                  # somebody's region overflowed: everybody retries with larger ones
        matrix = full[:, :W]
        plan = ExchangePlan(matrix, self.rank)
        side.plan = plan
        me = self.rank
        order = [(me + i) % W for i in range(W)]         # my own region first, then the peers behind me: no two ranks start on the same source
        plan = ExchangePlan(matrix, me, src_order=order)
        side.plan = plan
        side.piece_kw = okw
        side.pieces = [(side.send_pk.ptrs[s] + me * cap * okw, int(matrix[s, me])) for s in order]
        side.rid_pieces = [(side.send_pr.ptrs[s] + me * cap * 4, int(matrix[s, me])) for s in order]
        n_null = int(counts[W])
        side.null_keys = torch.as_tensor(_RawCudaArray(side.send_pk.local + W * cap * okw, n_null, "<i8" if okw == 8 else "<i4"), device=keys.device) if n_null else None
        side.null_rid = torch.as_tensor(_RawCudaArray(side.send_pr.local + W * cap * 4, n_null, "<i4"), device=keys.device) if n_null else None
        side.n_hot = 0
        if hot:
            side.n_hot = int(counts[W + 1])
            m = max(side.n_hot, 1)
            hk_ = torch.as_tensor(_RawCudaArray(side.send_pk.local + (W + 1) * cap * okw, m, "<i8" if okw == 8 else "<i4"), device=keys.device)
            side.hot_keys = hk_ if okw == kw else (hk_.to(torch.int64) & 0xFFFFFFFF)
            side.hot_rid = torch.as_tensor(_RawCudaArray(side.send_pr.local + (W + 1) * cap * 4, m, "<i4"), device=keys.device)
        side.recv_keys_ptr = side.recv_rid_ptr = None
        return int(plan.recv_rows[me])

    # ---- output rows across ranks (late materialisation; rows.py, csrc/gather_remote.cuh) ---------------------------------
    def _origin_pieces(self, side: _Side, rowid_base: int = 0, copy: bool = False):
        """OriginPieces of the batch `side` received last: per source rank (in the order the local indices count them) the row
        count and where that piece's row ids are -- in the peers' send regions (mode "pull") or in my receive buffer.
        copy=True keeps private copies of the row ids (the build side: its regions are overwritten by the next build batch)."""
        from .rows import OriginPieces
        torch = self.torch
        o = OriginPieces()
        plan, me = side.plan, self.rank
        if self.mode == "pull":          # one piece per source rank, in the rotated order the partition pass pulls them
            spans = [(int(s), int(m), int(p)) for s, (p, m) in zip(plan.src_order, side.rid_pieces)]
        else:                            # one receive buffer, the sources' segments in plan.src_order
            base = side.rid_pieces[0][0]
            spans = [(int(s), int(plan.counts[s, me]), int(base) + int(plan.src_offset[k]) * 4) for k, s in enumerate(plan.src_order)]
        for s, m, p in spans:
            if copy and m:
                t = torch.as_tensor(_RawCudaArray(p, m, "<i4"), device=torch.device("cuda", torch.cuda.current_device())).clone()
                o.add(s, m, t.data_ptr(), 0, keep=t)
            else:
                o.add(s, m, p if m else 0, 0)
        return o

    def share_columns(self, columns: Sequence):
        """Collective. Makes this rank's columns of one side readable by every rank: each pyarrow array (fixed width or Utf8,
        NULLs allowed) or device tensor is copied into cudaIpc-shared device buffers and the handles are all-gathered.
        Returns one rows.SharedColumn per column (per-rank pointers). Keep the result alive while rows are gathered from it;
        release it with unshare_columns (collective)."""
        from .rows import SharedColumn, arrow_to_device
        from .join import T_UTF8
        torch, dist = self.torch, self.dist
        dev = torch.device("cuda", torch.cuda.current_device())
        out = []
        for col in columns:
            if _is_torch_tensor(col):
                d_ = {"type": col.element_size(), "values": col.contiguous().view(torch.uint8), "validity": None, "data": None}
            else:
                d_ = arrow_to_device(torch, col, dev)
            bufs = {}
            # every rank must open every buffer: a rank without NULLs still allocates a (full) bitmap so that the collective is uniform
            n_rows = int(d_.get("n", 0)) if not _is_torch_tensor(col) else int(col.shape[0])
            if d_["validity"] is None:
                d_["validity"] = torch.full(((n_rows + 7) // 8 + 1,), 0xFF, dtype=torch.uint8, device=dev)
            for name in ("values", "validity", "data"):
                t = d_[name]
                if t is None:
                    continue
                nbytes = int(t.numel() * t.element_size())
                pb = _PeerBuffer(self.xjoin, dist, nbytes + 64)
                if nbytes:
                    torch.as_tensor(_RawCudaArray(pb.local, nbytes, "|u1"), device=dev).copy_(t.contiguous().view(torch.uint8).reshape(-1))
                bufs[name] = pb
            torch.cuda.synchronize()
            dist.barrier(group=self.group)
            out.append(SharedColumn(d_["type"], bufs["values"].ptrs, bufs["validity"].ptrs,
                                    bufs["data"].ptrs if "data" in bufs else None, keep=bufs))
        return out

    def unshare_columns(self, shared: Sequence):
        for c in shared:
            for pb in (c.keep or {}).values():
                pb.close(self.dist)
            c.keep = None

    def build_rows(self, idx, n: int, idx_width: int = 8):
        """rows.RowGatherer over LOCAL build indices (a probe result's build_idx, the finaliser's): gather(shared column) reads
        the values from the ranks that own the rows."""
        from .rows import RowGatherer, OriginPieces
        if not self.track_origin:
            raise B200JoinError(-5, "DistributedJoin(track_origin=True) is needed to map build rows back to their sources")
        o = OriginPieces()
        for part in self.build_origin:
            for k in range(len(part)):
                o.add(part.src[k], part.rows[k], part.rid_ptr[k], part.rid_base[k])
        return RowGatherer(self.join, self.torch, idx, n, idx_width, o)

    def probe_rows(self, idx, n: int, side: Optional[_Side] = None):
        """rows.RowGatherer over LOCAL probe indices of the batch `side` (default: the last one) received."""
        from .rows import RowGatherer
        side = side if side is not None else self.probe
        return RowGatherer(self.join, self.torch, idx, n, 4, self._origin_pieces(side))

    def null_build_rows(self):
        """Row ids (int32 device tensor) of this rank's build rows whose key is NULL (never sent; needs track_origin): Left /
        Full / LeftAnti joins emit them here, with a NULL probe side."""
        torch = self.torch
        if not self.build_null_rid:
            return torch.empty(0, dtype=torch.int32, device=torch.device("cuda", torch.cuda.current_device()))
        return torch.cat(self.build_null_rid)

    def null_probe_rows(self, side: Optional[_Side] = None):
        """Row ids (int32 device tensor) of this rank's probe rows whose key is NULL: they were never sent. Right / Full /
        RightAnti joins emit them here, with a NULL build side."""
        side = side if side is not None else self.probe
        return side.null_rid

    def received(self, side: _Side, what: str = "keys"):
        """The rows this rank received for one side as ONE local tensor (copies: tests, verification, late materialisation)."""
        torch = self.torch
        pieces = side.pieces if what == "keys" else side.rid_pieces
        pkw = getattr(side, "piece_kw", self.kw)
        ts = "<i4" if (what != "keys" or pkw == 4) else "<i8"
        dev = torch.device("cuda", torch.cuda.current_device())
        parts = [torch.as_tensor(_RawCudaArray(p, m, ts)).to(dev) for p, m in pieces if m]
        if not parts:
            return torch.empty(0, dtype=torch.int32 if ts == "<i4" else torch.int64, device=dev)
        out = torch.cat(parts)
        if what == "keys" and pkw != self.kw:           # 32-bit pieces of 8-byte keys: zero-extended
            out = out.to(torch.int64) & 0xFFFFFFFF
        return out

    def _local_probe(self, partition: int, side: _Side, want: int):
        pieces = [(p, m) for p, m in side.pieces]
        pkw = getattr(side, "piece_kw", self.kw)
        if len(pieces) == 1 and pkw == self.kw:
            p, m = pieces[0]
            return self.join.probe_batch(partition, [_dev_column(p, m, self.kw, keep=(side.keys,))], want=want)
        return self.join.probe_pieces(partition, pieces, want=want, key_width=pkw if pkw != self.kw else 0)

    # ---- the HashJoin call shape --------------------------------------------------------------------------------
    def _timed(self, which: str, fn):
        torch = self.torch
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(self.stream)
        out = fn()
        e1.record(self.stream)
        e1.synchronize()
        self.ms_exchange[which] += e0.elapsed_time(e1)
        return out

    def reset(self):
        self.join.reset()
        self.build_rows_local = 0
        self.build_rows_sent = 0
        self.build_origin = []
        self.build_null_rid = []
        if self.hot_join is not None:
            self.hot_join.close()
        self.hot = None
        self.hot_join = None
        for jh in (self.join, self.xjoin):
            shuffle_set_hot(jh, None)

    def build_push(self, partition: int, keys: Sequence):
        """Exchange this rank's build rows, then append what this rank received to its local build side."""
        k = keys[0]
        base = self.build_rows_sent                 # row ids count this rank's build rows over all its batches
        n_recv = self._timed("build", lambda: self._exchange(self.build, k, rowid_base=base))
        self.build_rows_sent += int(k.shape[0])
        self.build_rows_local += n_recv
        if self.track_origin:
            self.build_origin.append(self._origin_pieces(self.build, copy=True))
            if self.build.null_rid is not None and int(self.build.null_rid.shape[0]):
                self.build_null_rid.append(self.build.null_rid.clone())      # the region is overwritten by the next build batch
        for ptr, m in self.build.pieces:          # mode "pull": one piece per source rank, copied out of the peers' send regions
            if m:
                self.join.build_push(partition, [_dev_column(ptr, m, self.kw, keep=(self.build.keys,))])

    def build_finish(self, partition: int):
        self.join.build_finish(partition)

    # ---- heavy hitters (SURVEY §8e "skew") ------------------------------------------------------------------------
    # Left-family joins: the visited marks of the broadcast (hot) build rows live on every rank; probe_finish OR-reduces them
    # and the owner alone emits / withholds those rows ("OR-reduce those few bits").
    HOT_JOIN_TYPES = ("inner", "right", "right_semi", "right_anti", "left", "full", "left_semi", "left_anti")
    LEFT_FAMILY = ("left", "full", "left_semi", "left_anti")

    def enable_hot_routing(self, sample_keys, hot_fraction: Optional[float] = None, sample_rows: int = 1 << 20, max_hot: int = 256):
        """Collective, after build_finish. Finds the heavy-hitter probe keys from a sample of this rank's probe keys, agrees
        on their union over all ranks, BROADCASTS the build rows of those keys (every rank looks the hot keys up in its own
        table; the matches are all-gathered and loaded into a small local join), and tells the exchange kernels to keep rows
        with these keys where they are. A key holding the fraction f of all probe rows would otherwise send f * N rows to one
        GPU whose fair share is N / G. Returns the number of hot keys (0: routing stays off)."""
        torch, dist = self.torch, self.dist
        if self.join_type not in self.HOT_JOIN_TYPES:
            raise ValueError(f"heavy-hitter routing supports {self.HOT_JOIN_TYPES}, not {self.join_type}")
        W = self.world
        if hot_fraction is None:
            hot_fraction = 1.0 / (32.0 * W)
        n = int(sample_keys.shape[0])
        stride = max(1, n // sample_rows)
        samp = sample_keys[::stride][:sample_rows]
        uniq, cnt = torch.unique(samp, return_counts=True)
        sel = cnt.to(torch.float64) >= hot_fraction * float(samp.shape[0])
        uniq, cnt = uniq[sel], cnt[sel]
        if int(uniq.shape[0]) > max_hot:
            top = torch.topk(cnt, max_hot).indices
            uniq = uniq[top]
        mine = torch.zeros(max_hot, dtype=torch.int64, device=sample_keys.device)          # 0 = padding (never hot)
        mine[: int(uniq.shape[0])] = uniq.to(torch.int64) if self.kw == 8 else (uniq.to(torch.int64) & 0xFFFFFFFF)
        allk = torch.empty(W * max_hot, dtype=torch.int64, device=sample_keys.device)
        dist.all_gather_into_tensor(allk, mine, group=self.group)
        hot = torch.unique(allk)
        hot = hot[hot != 0]
        self.hot = None
        self.hot_join = None
        if int(hot.shape[0]) == 0:
            for jh in (self.join, self.xjoin):
                shuffle_set_hot(jh, None)
            return 0
        # the owners' build rows of the hot keys -> everyone
        hk_dev = hot if self.kw == 8 else hot.to(torch.int32)
        from .join import OUT_MATCHES_ONLY
        res = self.join.probe_batch(0, [hk_dev], want=OUT_PAIRS | OUT_MATCHES_ONLY)
        b_, p_ = res.device_arrays()
        if res.n:
            bi = torch.as_tensor(b_, device=hot.device).cpu().numpy().astype(np.int64)
            pi = torch.as_tensor(p_, device=hot.device).cpu().numpy().astype(np.int64) & 0xFFFFFFFF
        else:
            bi = pi = np.zeros(0, dtype=np.int64)
        res.release()
        hot_np = hot.cpu().numpy()
        parts: List = [None] * W
        dist.all_gather_object(parts, (hot_np[pi], bi), group=self.group)
        keys_all = np.concatenate([p[0] for p in parts]).astype(np.int64)
        self.hot_owner = np.concatenate([np.full(len(p[0]), r, dtype=np.int32) for r, p in enumerate(parts)])
        self.hot_bidx = np.concatenate([p[1] for p in parts]).astype(np.int64)
        self.hot_build_keys = torch.from_numpy(keys_all).to(hot.device)
        if self.kw == 4:
            self.hot_build_keys = self.hot_build_keys.to(torch.int32)
        self.hot_join = HashJoin(self.join_type, key_widths=(self.kw,), stream=self.stream.cuda_stream,
                                 device=torch.cuda.current_device())
        self.hot_join.build_push(0, [self.hot_build_keys])
        self.hot_join.build_finish(0)
        self.hot = hot_np.astype(np.uint64)
        return int(self.hot.shape[0])

    def _probe_hot(self, side: _Side, res, want: int):
        """Rows of hot keys never left this rank: join them against the broadcast build rows. res.hot = PairResult whose
        build_idx indexes hot_build_keys / hot_owner / hot_bidx (owner rank and its local build row) and whose probe_idx
        indexes side.hot_rid (the row's index in this rank's probe batch)."""
        res.hot = None
        res.n_hot = 0
        if self.hot_join is None:
            return res
        hk = side.hot_keys[: side.n_hot]
        res.hot = self.hot_join.probe_batch(0, [hk], want=want)
        res.n_hot = res.hot.n
        res.hot_rid = side.hot_rid[: side.n_hot]
        return res

    def probe_batch(self, partition: int, keys: Sequence, want: int = OUT_PAIRS):
        """Exchange one probe batch, probe the received rows locally. The result's indices are LOCAL: build_idx counts
        the build rows this rank received (in arrival order), probe_idx the rows of the received probe batch; map them to
        (source rank, source row) with `origin_build` / `origin_probe`."""
        k = keys[0]
        n_recv = self._timed("probe", lambda: self._exchange(self.probe, k, rowid_base=0))
        res = self._local_probe(partition, self.probe, want)
        res.exchange_plan = self.probe.plan
        return self._probe_hot(self.probe, res, want)

    def probe_exchange_begin(self, keys: Sequence, side_index: int = 0):
        """Starts the exchange of one probe batch on the exchange handle / stream from a helper thread and returns a token for
        probe_batch_received. Meant to run beside build_finish: the NVLink exchange of the probe side then overlaps the local
        table build (which touches neither the probe keys nor the receive buffers). Collective: every rank calls it."""
        import threading
        torch = self.torch
        dev = torch.cuda.current_device()
        side = self.probe_sides[side_index]
        box = {}

        def run():
            try:
                torch.cuda.set_device(dev)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(self.xstream)
                box["n"] = self._exchange(side, keys[0], rowid_base=0, overlapped=True)
                e1.record(self.xstream)
                e1.synchronize()
                self.ms_exchange["probe"] += e0.elapsed_time(e1)
            except BaseException as exc:
                box["n"] = exc
        th = threading.Thread(target=run, daemon=True)
        th.start()
        return (th, box, side)

    def probe_batch_received(self, partition: int, token, want: int = OUT_PAIRS):
        """Joins the rows a probe_exchange_begin delivered (local probe on the join's own stream)."""
        th, box, side = token
        th.join()
        if isinstance(box["n"], BaseException):
            raise box["n"]
        self.probe = side
        res = self._local_probe(partition, side, want)
        res.exchange_plan = side.plan
        return self._probe_hot(side, res, want)

    def probe_chunks(self, partition: int, key_chunks: Sequence, want: int = OUT_PAIRS):
        """Generator over the results of a sequence of probe chunks (device key tensors), with the partition + exchange
        of chunk k+1 running on the exchange stream while chunk k is joined locally ("chunked and
        overlapped"). Every rank must iterate over the same number of chunks. The exchange of chunk k+2 overwrites the
        receive buffers of chunk k: it is started only after chunk k's local join has returned, and its count all-gather
        (every rank's exchange thread takes part) keeps any peer from storing into buffers a slower rank still reads."""
        import threading
        torch = self.torch
        chunks = list(key_chunks)
        dev = torch.cuda.current_device()
        box = {}

        def run_exchange(k):
            try:
                torch.cuda.set_device(dev)                      # the CUDA device is per thread
                side = self.probe_sides[k & 1]
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(self.xstream)
                n_recv = self._exchange(side, chunks[k], rowid_base=0, overlapped=True)
                e1.record(self.xstream)
                e1.synchronize()
                self.ms_exchange["probe"] += e0.elapsed_time(e1)
                box[k] = n_recv
            except BaseException as exc:                        # surfaces in the consuming thread
                box[k] = exc

        def start(k):
            th = threading.Thread(target=run_exchange, args=(k,), daemon=True)
            th.start()
            return th

        if not chunks:
            return
        th = start(0)
        for k in range(len(chunks)):
            th.join()
            got = box.pop(k)
            if isinstance(got, BaseException):
                raise got
            th = start(k + 1) if k + 1 < len(chunks) else None
            side = self.probe_sides[k & 1]
            self.probe = side
            res = self._local_probe(partition, side, want)
            res.exchange_plan = side.plan
            yield self._probe_hot(side, res, want)

    # ---- host batches (pinned numpy in, pinned index pairs out), copies pipelined around the exchange + join ------------
    def _host_state(self, chunk_rows: int, out_rows: int):
        """Device staging buffers (two chunks of keys), copy streams and the pinned result arrays; grown, never shrunk."""
        torch = self.torch
        hs = getattr(self, "_hs", None)
        if hs is None:
            dev = torch.cuda.current_device()
            hs = self._hs = {"in": torch.cuda.Stream(device=dev), "out": torch.cuda.Stream(device=dev), "chunk": 0, "cap": 0}
        if chunk_rows > hs["chunk"]:
            hs["dbuf"] = [torch.empty(chunk_rows, dtype=self._key_dtype, device="cuda") for _ in range(2)]
            hs["chunk"] = chunk_rows
        if out_rows > hs["cap"]:
            ob = torch.empty(out_rows, dtype=torch.int64).pin_memory()
            op = torch.empty(out_rows, dtype=torch.int32).pin_memory()
            if hs["cap"]:
                hs["out"].synchronize()
                ob[: hs["cap"]].copy_(hs["out_b"])
                op[: hs["cap"]].copy_(hs["out_p"])
            hs["out_b"], hs["out_p"], hs["cap"] = ob, op, out_rows
        return hs

    def build_push_host(self, partition: int, host_keys: np.ndarray):
        """build_push for a host (ideally pinned) key array: one H2D copy on the join's stream, then the exchange."""
        torch = self.torch
        with torch.cuda.stream(self.stream):
            k = torch.from_numpy(host_keys).to("cuda", non_blocking=True)
            self.build_push(partition, [k])

    def probe_host_chunks(self, partition: int, host_keys: np.ndarray, chunk_rows: int = 1 << 27, want: int = OUT_PAIRS):
        """Generator over the chunks of ONE host probe batch (a pinned numpy key array): the H2D copy of chunk k+1 and
        the D2H copy of chunk k-1's pairs run on their own streams while chunk k is exchanged and joined, so that the
        PCIe link works in both directions for the whole batch (the multi-GPU counterpart of the C ABI's pipelined host
        probe). Yields, for every chunk in order, the probe result with `host_build_idx` / `host_probe_idx` set to numpy
        views of this rank's pinned result arrays (valid until the next call of this method) and `side` = the receive
        buffers the chunk's LOCAL indices refer to (`origin_probe(idx, side=res.side)`; they survive one more chunk).
        Every rank must call with the same number of chunks."""
        torch = self.torch
        hk = torch.from_numpy(host_keys)
        n = int(hk.shape[0])
        chunk_rows = max(1, min(chunk_rows, n))
        hs = self._host_state(chunk_rows, int(n * 1.05) + 4096)
        s_in, s_out = hs["in"], hs["out"]
        bounds = [(lo, min(chunk_rows, n - lo)) for lo in range(0, n, chunk_rows)]
        ev_in = [torch.cuda.Event(), torch.cuda.Event()]

        def stage(k):
            lo, m = bounds[k]
            with torch.cuda.stream(s_in):
                hs["dbuf"][k & 1][:m].copy_(hk[lo:lo + m], non_blocking=True)
                ev_in[k & 1].record(s_in)

        def finish(item):
            res, ev, off = item
            ev.synchronize()
            if res.build_idx_width == 4:                 # B200J_OUT_BUILD_U32: 4-byte local build indices (8 instead of 12 bytes per pair over PCIe)
                res.host_build_idx = hs["out_b"].view(torch.int32)[off:off + res.n].numpy().view(np.uint32)
            else:
                res.host_build_idx = hs["out_b"][off:off + res.n].numpy().view(np.uint64)
            res.host_probe_idx = hs["out_p"][off:off + res.n].numpy().view(np.uint32)
            res.release()
            return res

        if not bounds:
            return
        stage(0)
        pending, off = None, 0
        for k, (lo, m) in enumerate(bounds):
            # probe_batch(k-1) returned with its result counted: the staging buffer of chunk k-1 is consumed, chunk k+1 may land in it
            if k + 1 < len(bounds):
                stage(k + 1)
            self.stream.wait_event(ev_in[k & 1])
            self.probe = self.probe_sides[k & 1]
            with torch.cuda.stream(self.stream):
                res = self.probe_batch(partition, [hs["dbuf"][k & 1][:m]], want=want)
            res.side, res.chunk_lo = self.probe, lo
            if off + res.n > hs["cap"]:
                hs = self._host_state(chunk_rows, int((off + res.n) * 1.5) + 4096)
            done = torch.cuda.Event()
            done.record(self.stream)
            ev_out = torch.cuda.Event()
            with torch.cuda.stream(s_out):
                s_out.wait_event(done)
                if res.n:
                    b_, p_ = res.device_arrays()
                    if res.build_idx_width == 4:
                        hs["out_b"].view(torch.int32)[off:off + res.n].copy_(torch.as_tensor(b_, device="cuda").view(torch.int32), non_blocking=True)
                    else:
                        hs["out_b"][off:off + res.n].copy_(torch.as_tensor(b_, device="cuda").view(torch.int64), non_blocking=True)
                    hs["out_p"][off:off + res.n].copy_(torch.as_tensor(p_, device="cuda").view(torch.int32), non_blocking=True)
                ev_out.record(s_out)
            if pending is not None:
                yield finish(pending)
            pending = (res, ev_out, off)
            off += res.n
        yield finish(pending)

    def probe_finish(self, partition: int, want: int = OUT_PAIRS):
        """The finaliser batch of this rank's build rows (Left / Full / LeftSemi / LeftAnti; empty otherwise).
        With heavy-hitter routing the build rows of hot keys were probed on EVERY rank (against the broadcast copy), never in
        their owner's table: collective then. Every rank turns its hot join's finaliser into one "matched here" flag per
        broadcast row, the flags are OR-reduced over the ranks, and each owner drops its hot rows from its own finaliser
        batch and puts back the ones the join type wants (unmatched everywhere: Left / Full / LeftAnti; matched anywhere:
        LeftSemi). Returns host arrays in that case (`HostPairs`)."""
        if self.hot_join is None or self.join_type not in self.LEFT_FAMILY:
            return self.join.probe_finish(partition, want=want)
        torch, dist = self.torch, self.dist
        from .join import OUT_HOST
        main = self.join.probe_finish(partition, want=OUT_PAIRS | OUT_HOST)
        mb, _ = main.numpy()
        main.release()
        hot = self.hot_join.probe_finish(0, want=OUT_PAIRS | OUT_HOST)
        hb, _ = hot.numpy()
        hot.release()
        n_hot_rows = int(self.hot_bidx.shape[0])
        if self.join_type == "left_semi":
            flag = np.zeros(n_hot_rows, dtype=np.int32)             # finaliser = matched rows
            flag[hb.astype(np.int64)] = 1
        else:
            flag = np.ones(n_hot_rows, dtype=np.int32)              # finaliser = unmatched rows
            flag[hb.astype(np.int64)] = 0
        t = torch.from_numpy(flag).to("cuda")
        if n_hot_rows:
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        matched = t.cpu().numpy().astype(bool)
        mine = self.hot_owner == self.rank
        my_bidx = self.hot_bidx[mine].astype(np.uint64)
        keep = ~np.isin(mb, my_bidx)
        back = my_bidx[matched[mine]] if self.join_type == "left_semi" else my_bidx[~matched[mine]]
        b = np.concatenate([mb[keep], back]).astype(np.uint64)
        from .join import NULL_PROBE
        return HostPairs(b, np.full(b.shape[0], NULL_PROBE, dtype=np.uint32))

    def _origin(self, side: _Side, idx: np.ndarray):
        """(source rank, source row) of received rows `idx` of one side (host arrays; for tests and late materialisation)."""
        rid = self.received(side, "rid").cpu().numpy().view(np.uint32)
        src_rank, _ = side.plan.origin_of(np.asarray(idx, dtype=np.int64))
        return src_rank, rid[np.asarray(idx, dtype=np.int64)]

    def origin_build(self, build_idx):
        return self._origin(self.build, build_idx)

    def origin_probe(self, probe_idx, side: Optional[_Side] = None):
        return self._origin(side if side is not None else self.probe, probe_idx)

    def stats(self) -> dict:
        st = self.join.stats()
        st["ms_exchange_build"] = self.ms_exchange["build"]
        st["ms_exchange_probe"] = self.ms_exchange["probe"]
        return st

    def reset_stats(self):
        self.join.reset_stats()
        self.ms_exchange = {"build": 0.0, "probe": 0.0}

    def close(self):
        if self.mode in ("p2p", "ce", "pull"):
            for side in [self.build] + self.probe_sides:
                for b in (side.keys, side.rid):
                    if b is not None:
                        b.close(self.dist)
                side.keys = side.rid = None
                for b in (getattr(side, "send_pk", None), getattr(side, "send_pr", None)):
                    if b is not None:
                        b.close(self.dist)
                side.send_pk = side.send_pr = None
        if self.hot_join is not None:
            self.hot_join.close()
        self.xjoin.close()
        self.join.close()


def _is_torch_tensor(x) -> bool:
    return type(x).__module__.startswith("torch") and hasattr(x, "data_ptr")


class DistributedKeyJoin:
    """Join keyed on SEVERAL columns and / or Utf8 columns over the ranks ("composite / Utf8 keys through the
    exchange"). The exchange and the local joins work on the ROW HASH of the key columns (create_hashes, b200j_hash_keys:
    what the reference's table is keyed on, src/shared/shared.rs:11-16) as an 8-byte key; what the local join finds are
    CANDIDATE pairs (B200J_OUT_MATCHES_ONLY). Their key columns are then read from the ranks that own the rows and compared
    (equal_rows_arr, datafusion_private.rs:40-80: rows.equal_key_rows), and the surviving pairs go through the join-type
    algebra of b200j_apply_pairs -- the same round trip a JoinFilter takes. Rows with a NULL key travel like any other row
    and lose their candidates in the check, so Right / Full joins emit them from the rank that received them.
    Key columns are pyarrow arrays; results are LOCAL index pairs: materialise them with .dj.build_rows / .dj.probe_rows."""

    def __init__(self, join_type="inner", key_types: Sequence[int] = (8,), **kwargs):
        import torch
        self.torch = torch
        kwargs.setdefault("mode", "pull")
        self.dj = DistributedJoin(join_type, key_width=8, track_origin=True, narrow_keys=False, **kwargs)
        self.hasher = HashJoin("inner", key_widths=tuple(key_types), device=torch.cuda.current_device())
        self._build_batches: List[list] = []
        self.build_key_cols = None
        self.probe_key_cols = None

    def _hashes(self, key_arrays):
        h = self.hasher.hash_keys(list(key_arrays)).view(np.int64)
        return self.torch.from_numpy(h.copy()).to(self.torch.device("cuda", self.torch.cuda.current_device()))

    def build_push(self, partition: int, key_arrays: Sequence):
        self._build_batches.append(list(key_arrays))
        self.dj.build_push(partition, [self._hashes(key_arrays)])

    def build_finish(self, partition: int):
        import pyarrow as pa
        n_cols = len(self._build_batches[0])
        cols = [pa.concat_arrays([b[c].combine_chunks() if isinstance(b[c], pa.ChunkedArray) else b[c] for b in self._build_batches])
                for c in range(n_cols)]
        self.build_key_cols = self.dj.share_columns(cols)
        self.dj.build_finish(partition)

    def probe_batch(self, partition: int, key_arrays: Sequence, want: int = OUT_PAIRS):
        from .join import OUT_MATCHES_ONLY
        from .rows import equal_key_rows
        torch, dj = self.torch, self.dj
        if self.probe_key_cols is not None:
            dj.unshare_columns(self.probe_key_cols)
        self.probe_key_cols = dj.share_columns(list(key_arrays))
        cand = dj.probe_batch(partition, [self._hashes(key_arrays)], want=OUT_PAIRS | OUT_MATCHES_ONLY)
        n_recv = int(dj.probe.plan.recv_rows[dj.rank])
        dev = torch.device("cuda", torch.cuda.current_device())
        if cand.n:
            cb, cp = cand.device_arrays()
            cb_t, cp_t = torch.as_tensor(cb, device=dev), torch.as_tensor(cp, device=dev)
            keep = equal_key_rows(dj.join, torch, dj.build_rows(cb_t, cand.n, cand.build_idx_width), dj.probe_rows(cp_t, cand.n),
                                  self.build_key_cols, self.probe_key_cols)
            kb, kp = cb_t[keep].contiguous(), cp_t[keep].contiguous()
        else:
            kb = torch.empty(0, dtype=torch.int64, device=dev)
            kp = torch.empty(0, dtype=torch.int32, device=dev)
        torch.cuda.current_stream().synchronize()
        res = dj.join.apply_pairs_device(partition, n_recv, kb.data_ptr(), kp.data_ptr(), int(kb.shape[0]), want=want)
        cand.release()
        res.exchange_plan = dj.probe.plan
        return res

    def probe_finish(self, partition: int, want: int = OUT_PAIRS):
        return self.dj.probe_finish(partition, want=want)

    def close(self):
        for cols in (self.probe_key_cols, self.build_key_cols):
            if cols is not None:
                self.dj.unshare_columns(cols)
        self.hasher.close()
        self.dj.close()


class HostPairs:
    """A finished result held in host arrays (same reading interface as PairResult)."""

    def __init__(self, build_idx: np.ndarray, probe_idx: np.ndarray):
        self.build_idx, self.probe_idx = build_idx, probe_idx
        self.n = int(build_idx.shape[0])
        self.hot = None
        self.n_hot = 0

    def numpy(self):
        return self.build_idx, self.probe_idx

    def release(self):
        pass


class _RawCudaArray:
    def __init__(self, ptr: int, n: int, typestr: str):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr or 0, False), "version": 3, "strides": None}
