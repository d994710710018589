// gather.cuh — output gather (take_multiple_record_batch, src/shared/shared.rs:83-92) and the build-store helpers.
#pragma once
#include "kernels.cuh"
namespace b200j {
// Utf8 concat: dst[i] = src[i] - first + base  (Arrow 32-bit offsets of one batch -> 64-bit offsets of the build store)
__global__ void __launch_bounds__(256) rebase_offsets_kernel(const int32_t *__restrict__ src, int64_t n, int64_t first, int64_t base,
                                                             int64_t *__restrict__ dst) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        dst[i] = (int64_t)src[i] - first + base;
}
}  // namespace b200j

// ---------------------------------------------------------------------------------------------------------------------
// a6: output gather — take_multiple_record_batch (src/shared/shared.rs:83-92) = arrow::compute::take of every build
// column at build_idx and every probe column at probe_idx. A NULL index (~0) yields a NULL output row, as does a NULL
// source row. One thread per output row; a warp covers 32 consecutive rows, so its ballot IS the output validity word.
// Bound: HBM — per output row the column width read (random on the build side, near-sequential on the probe side)
// plus the width written.
// ---------------------------------------------------------------------------------------------------------------------
namespace b200j {

template <typename IdxT>
__device__ __forceinline__ bool idx_is_null(IdxT v) { return v == (IdxT)~(IdxT)0; }

struct GatherSrc {
    const void    *values;       // fixed width: values; Utf8: data bytes
    const void    *offsets;      // Utf8: int64 (build store) or int32 (probe batch) offsets
    const uint8_t *validity;     // LSB-first bitmap or null
    int64_t        bit_offset;
    int32_t        offsets64;    // 1: offsets are int64
    int32_t        _pad;
};

template <typename T, typename IdxT>
__global__ void __launch_bounds__(256) gather_fixed_kernel(GatherSrc src, const IdxT *__restrict__ idx, int64_t n, T *__restrict__ dst,
                                                           uint32_t *__restrict__ dst_valid, unsigned long long *null_count) {
    const int lane = threadIdx.x & 31;
    uint32_t nulls = 0;
    // rows are handled in warp-aligned groups of 32 so every validity word has exactly one writer
    const int64_t n_round = (n + 31) & ~(int64_t)31;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += (int64_t)gridDim.x * blockDim.x) {
        bool valid = false;
        T v{};
        if (i < n) {
            const IdxT j = idx ? idx[i] : (IdxT)~(IdxT)0;
            if (!idx_is_null(j) && src.values != nullptr && bit_valid(src.validity, src.bit_offset + (int64_t)j)) {
                v = reinterpret_cast<const T *>(src.values)[j];
                valid = true;
            }
            dst[i] = v;
            if (!valid) nulls++;
        }
        const uint32_t word = __ballot_sync(0xffffffffu, valid);
        if (lane == 0 && dst_valid) dst_valid[i >> 5] = word;
    }
    const unsigned long long z = warp_sum_u32(nulls);
    if (lane == 0 && z) atomicAdd(null_count, z);
}

// All fixed-width columns of one side in ONE launch (round 2). A warp takes 128 consecutive output rows, lane l the rows
// l, l + 32, l + 64, l + 96 of them: the index loads and every column's stores are lane-contiguous, a ballot per item IS the
// output validity word, and each thread has 4 independent random reads per column in flight (the gather is bound by the
// latency and sector traffic of those reads: 32 B moved per value). The index array is read once for all columns.
constexpr int GF_MAX_COLS = 8;
constexpr int GF_ITEMS = 4;
struct GatherCol {
    const void    *values;       // null: the side has no rows / the finaliser's NULL probe side
    const uint8_t *validity;
    int64_t        bit_offset;
    void          *dst;
    uint32_t      *dst_valid;
    int32_t        width;        // 1, 2, 4, 8, 16
    int32_t        _pad;
};
struct GatherCols {
    GatherCol c[GF_MAX_COLS];
    int32_t   n;
    int32_t   _pad;
};

template <typename T>
__device__ __forceinline__ void gather_items(const GatherCol &gc, const int64_t (&j)[GF_ITEMS], const bool (&have)[GF_ITEMS], int64_t row0, int lane,
                                             int64_t n, uint32_t &nulls) {
    T v[GF_ITEMS];
    bool valid[GF_ITEMS];
#pragma unroll
    for (int q = 0; q < GF_ITEMS; q++) {
        valid[q] = have[q] && gc.values != nullptr && bit_valid(gc.validity, gc.bit_offset + j[q]);
        v[q] = T{};
        if (valid[q]) v[q] = reinterpret_cast<const T *>(gc.values)[j[q]];
    }
#pragma unroll
    for (int q = 0; q < GF_ITEMS; q++) {
        const int64_t i = row0 + q * 32 + lane;
        if (i < n) {
            reinterpret_cast<T *>(gc.dst)[i] = v[q];
            if (!valid[q]) nulls++;
        }
        const uint32_t word = __ballot_sync(0xffffffffu, valid[q]);
        if (lane == 0 && row0 + q * 32 < n) gc.dst_valid[(row0 + q * 32) >> 5] = word;
    }
}

template <typename IdxT>
__global__ void __launch_bounds__(256) gather_fixed_multi_kernel(GatherCols cols, const IdxT *__restrict__ idx, int64_t n,
                                                                 unsigned long long *null_counts /* [cols.n] */) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    uint32_t nulls[GF_MAX_COLS];
#pragma unroll
    for (int c = 0; c < GF_MAX_COLS; c++) nulls[c] = 0;
    for (int64_t row0 = (((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5) * (32 * GF_ITEMS); row0 < n; row0 += warps * (32 * GF_ITEMS)) {
        int64_t j[GF_ITEMS];
        bool have[GF_ITEMS];
#pragma unroll
        for (int q = 0; q < GF_ITEMS; q++) {
            const int64_t i = row0 + q * 32 + lane;
            const IdxT x = (idx != nullptr && i < n) ? __ldcs(idx + i) : (IdxT)~(IdxT)0;
            have[q] = !idx_is_null(x);
            j[q] = have[q] ? (int64_t)x : 0;
        }
#pragma unroll
        for (int c = 0; c < GF_MAX_COLS; c++) {
            if (c >= cols.n) break;
            const GatherCol &gc = cols.c[c];
            switch (gc.width) {
                case 1: gather_items<uint8_t>(gc, j, have, row0, lane, n, nulls[c]); break;
                case 2: gather_items<uint16_t>(gc, j, have, row0, lane, n, nulls[c]); break;
                case 4: gather_items<uint32_t>(gc, j, have, row0, lane, n, nulls[c]); break;
                case 8: gather_items<unsigned long long>(gc, j, have, row0, lane, n, nulls[c]); break;
                default: gather_items<ulonglong2>(gc, j, have, row0, lane, n, nulls[c]); break;
            }
        }
    }
#pragma unroll
    for (int c = 0; c < GF_MAX_COLS; c++) {
        if (c >= cols.n) break;
        const unsigned long long z = warp_sum_u32(nulls[c]);
        if (lane == 0 && z) atomicAdd(null_counts + c, z);
    }
}

// The build KEY column of an equi-join's output rows, without touching the build store: a matched pair's build key IS its probe
// key (bit for bit: the table compares key values), and the probe key column is read at probe_idx -- near-sequential for the
// dense probe's output -- instead of a random read into the whole build store per row. A row whose probe index is NULL (the
// finaliser's rows, should a caller pass them with a batch) still reads the build store. Same 128-rows-per-warp shape as above.
template <typename T, typename IdxT>
__global__ void __launch_bounds__(256) gather_key_via_probe_kernel(const IdxT *__restrict__ bidx, const uint32_t *__restrict__ pidx, int64_t n,
                                                                   const T *__restrict__ build_key, const uint8_t *__restrict__ build_validity,
                                                                   const T *__restrict__ probe_key, T *__restrict__ dst, uint32_t *__restrict__ dst_valid,
                                                                   unsigned long long *null_count) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    uint32_t nulls = 0;
    for (int64_t row0 = (((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5) * (32 * GF_ITEMS); row0 < n; row0 += warps * (32 * GF_ITEMS)) {
        T v[GF_ITEMS];
        bool valid[GF_ITEMS];
#pragma unroll
        for (int q = 0; q < GF_ITEMS; q++) {
            const int64_t i = row0 + q * 32 + lane;
            v[q] = T{};
            valid[q] = false;
            if (i < n) {
                const IdxT b = __ldcs(bidx + i);
                const uint32_t p = __ldcs(pidx + i);
                if (!idx_is_null(b)) {
                    if (p != 0xffffffffu) {
                        v[q] = probe_key[p];
                        valid[q] = true;
                    } else if (build_key != nullptr && bit_valid(build_validity, (int64_t)b)) {
                        v[q] = build_key[b];
                        valid[q] = true;
                    }
                }
            }
        }
#pragma unroll
        for (int q = 0; q < GF_ITEMS; q++) {
            const int64_t i = row0 + q * 32 + lane;
            if (i < n) {
                dst[i] = v[q];
                if (!valid[q]) nulls++;
            }
            const uint32_t word = __ballot_sync(0xffffffffu, valid[q]);
            if (lane == 0 && row0 + q * 32 < n) dst_valid[(row0 + q * 32) >> 5] = word;
        }
    }
    const unsigned long long z = warp_sum_u32(nulls);
    if (lane == 0 && z) atomicAdd(null_count, z);
}

// Both KEY columns of an equi-join's output rows in ONE pass over the index pairs: the probe key column (probe_key[probe_idx])
// and the build key column, which for a matched pair is the same value (see gather_key_via_probe_kernel). One read of the
// pairs, one random read per row, two coalesced writes -- instead of two launches that each read the pairs and the probe key.
template <typename T, typename IdxT>
__global__ void __launch_bounds__(256) gather_keys_both_kernel(const IdxT *__restrict__ bidx, const uint32_t *__restrict__ pidx, int64_t n,
                                                               const T *__restrict__ build_key, const uint8_t *__restrict__ build_validity,
                                                               const T *__restrict__ probe_key, const uint8_t *__restrict__ probe_validity,
                                                               int64_t probe_bit_offset, T *__restrict__ dst_b, uint32_t *__restrict__ dst_b_valid,
                                                               T *__restrict__ dst_p, uint32_t *__restrict__ dst_p_valid,
                                                               unsigned long long *null_b, unsigned long long *null_p) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    uint32_t nulls_b = 0, nulls_p = 0;
    for (int64_t row0 = (((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5) * (32 * GF_ITEMS); row0 < n; row0 += warps * (32 * GF_ITEMS)) {
        T vb[GF_ITEMS], vp[GF_ITEMS];
        bool okb[GF_ITEMS], okp[GF_ITEMS];
#pragma unroll
        for (int q = 0; q < GF_ITEMS; q++) {
            const int64_t i = row0 + q * 32 + lane;
            vb[q] = T{};
            vp[q] = T{};
            okb[q] = okp[q] = false;
            if (i < n) {
                const IdxT b = __ldcs(bidx + i);
                const uint32_t p = __ldcs(pidx + i);
                if (p != 0xffffffffu) {
                    const T v = probe_key[p];
                    if (bit_valid(probe_validity, probe_bit_offset + (int64_t)p)) { vp[q] = v; okp[q] = true; }
                    if (!idx_is_null(b)) { vb[q] = v; okb[q] = true; }            // a matched pair: the keys are equal, so neither is NULL
                } else if (!idx_is_null(b) && build_key != nullptr && bit_valid(build_validity, (int64_t)b)) {
                    vb[q] = build_key[b];
                    okb[q] = true;
                }
            }
        }
#pragma unroll
        for (int q = 0; q < GF_ITEMS; q++) {
            const int64_t i = row0 + q * 32 + lane;
            if (i < n) {
                __stcs(dst_b + i, vb[q]);
                __stcs(dst_p + i, vp[q]);
                if (!okb[q]) nulls_b++;
                if (!okp[q]) nulls_p++;
            }
            const uint32_t wb = __ballot_sync(0xffffffffu, okb[q]), wp = __ballot_sync(0xffffffffu, okp[q]);
            if (lane == 0 && row0 + q * 32 < n) {
                dst_b_valid[(row0 + q * 32) >> 5] = wb;
                dst_p_valid[(row0 + q * 32) >> 5] = wp;
            }
        }
    }
    const unsigned long long zb = warp_sum_u32(nulls_b), zp = warp_sum_u32(nulls_p);
    if (lane == 0 && zb) atomicAdd(null_b, zb);
    if (lane == 0 && zp) atomicAdd(null_p, zp);
}

// ---------------------------------------------------------------------------------------------------------------------
// Build-side gather in PARTITION order (round 2). The dense probe behind the radix layout leaves its pairs in the order of
// the block-partitioned probe stream (radix_block.cuh): block after block, and inside a block partition after partition.
// Gathering the build columns in that order touches, per 32 K output rows, build rows of ALL partitions -- one random 32 B
// DRAM sector per row and column (measured: 26 ms per 1 B rows for one int64 column, the DRAM's random-sector rate). Walked
// the way the probe walks -- for p = 0 .. P: the segments (p, b) of all blocks -- the rows in flight reference the build rows
// of ONE partition: 1/P of the column's sectors (25 MB at C2), fetched from DRAM once and then served by the L2 for the
// ~10 output rows that share each of them. Reads of the index array and writes of the output stay coalesced (a segment is
// a contiguous run of output rows). Validity words straddle segments that different warps write at different times, so the
// bitmaps are pre-set to all-valid by the host and NULL rows clear their bits with atomicAnd (rare).
// ---------------------------------------------------------------------------------------------------------------------
struct GatherSegArgs {
    const uint32_t *seg;          // [n_blocks][P + 2] segment starts inside each block (radix_block.cuh)
    const uint32_t *block_base;   // [n_blocks] first output row of every block
    uint32_t        n_blocks;
    uint32_t        P;
    uint32_t        n_units;      // n_blocks * (P + 1): the last "partition" holds the NULL-key rows
    unsigned int   *next;         // claim counter, zero at launch
};
constexpr int GS_UNITS = 4;       // segments per warp claim

// Reads of the build columns: no L1 allocation (a sector is never reused by the same SM soon enough) and an L2 evict-last
// policy (the current partition's sectors are the only data with reuse; index and output streams are evict-first).
__device__ __forceinline__ uint64_t l2_evict_last_policy() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
template <typename T>
__device__ __forceinline__ T ld_gather_keep(const T *p, uint64_t) { return __ldcg(p); }
template <>
__device__ __forceinline__ unsigned long long ld_gather_keep<unsigned long long>(const unsigned long long *p, uint64_t pol) {
    unsigned long long v;
    asm volatile("ld.global.L1::no_allocate.L2::cache_hint.u64 %0, [%1], %2;" : "=l"(v) : "l"(p), "l"(pol));
    return v;
}
template <>
__device__ __forceinline__ uint32_t ld_gather_keep<uint32_t>(const uint32_t *p, uint64_t pol) {
    uint32_t v;
    asm volatile("ld.global.L1::no_allocate.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(pol));
    return v;
}

template <typename T>
__device__ __forceinline__ void gather_seg_items(const GatherCol &gc, const int64_t (&j)[GF_ITEMS], const bool (&have)[GF_ITEMS],
                                                 const bool (&inr)[GF_ITEMS], int64_t row0, int lane, uint32_t &nulls, uint64_t pol) {
    T v[GF_ITEMS];
    bool valid[GF_ITEMS];
#pragma unroll
    for (int q = 0; q < GF_ITEMS; q++) {
        valid[q] = have[q] && gc.values != nullptr && bit_valid(gc.validity, gc.bit_offset + j[q]);
        v[q] = T{};
        if (valid[q]) v[q] = ld_gather_keep(reinterpret_cast<const T *>(gc.values) + j[q], pol);
    }
#pragma unroll
    for (int q = 0; q < GF_ITEMS; q++) {
        const int64_t r0 = row0 + q * 32;
        if (inr[q]) {
            __stcs(reinterpret_cast<T *>(gc.dst) + r0 + lane, v[q]);
            if (!valid[q]) nulls++;
        }
        const uint32_t inv = __ballot_sync(0xffffffffu, inr[q] && !valid[q]);
        if (inv != 0u && lane == 0) {
            const uint32_t sh = (uint32_t)(r0 & 31);
            const uint32_t lo = inv << sh;
            if (lo) atomicAnd(gc.dst_valid + (r0 >> 5), ~lo);
            if (sh) {
                const uint32_t hi = inv >> (32u - sh);
                if (hi) atomicAnd(gc.dst_valid + (r0 >> 5) + 1, ~hi);
            }
        }
    }
}

template <typename IdxT>
__global__ void __launch_bounds__(256) gather_seg_multi_kernel(GatherCols cols, const IdxT *__restrict__ idx, int64_t n, GatherSegArgs sg,
                                                               unsigned long long *null_counts /* [cols.n] */) {
    constexpr uint32_t FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const uint64_t pol = l2_evict_last_policy();
    uint32_t nulls[GF_MAX_COLS];
#pragma unroll
    for (int c = 0; c < GF_MAX_COLS; c++) nulls[c] = 0;
    // bits past the last row are zero in every bitmap this library hands out
    if (blockIdx.x == 0 && (int)threadIdx.x < cols.n && (n & 31) != 0)
        atomicAnd(cols.c[threadIdx.x].dst_valid + (n >> 5), (1u << (uint32_t)(n & 31)) - 1u);
    const uint32_t n_claims = (sg.n_units + GS_UNITS - 1) / GS_UNITS;
    uint32_t claim = 0;
    if (lane == 0) claim = atomicAdd(sg.next, 1u);
    for (;;) {
        const uint32_t cur = __shfl_sync(FULL, claim, 0);
        if (cur >= n_claims) break;
        if (lane == 0) claim = atomicAdd(sg.next, 1u);
        uint32_t my_start = 0, my_len = 0;
        {
            const uint32_t u = cur * GS_UNITS + (uint32_t)lane;
            if (lane < GS_UNITS && u < sg.n_units) {
                const uint32_t p = u / sg.n_blocks, b = u - p * sg.n_blocks;
                const uint32_t e0 = __ldg(sg.seg + (size_t)b * (sg.P + 2) + p), e1 = __ldg(sg.seg + (size_t)b * (sg.P + 2) + p + 1);
                my_start = __ldg(sg.block_base + b) + e0;
                my_len = e1 - e0;
            }
        }
#pragma unroll 1
        for (int s = 0; s < GS_UNITS; s++) {
            const uint32_t len = __shfl_sync(FULL, my_len, s);
            if (len == 0) continue;
            const int64_t start = (int64_t)__shfl_sync(FULL, my_start, s);
            for (uint32_t i0 = 0; i0 < len; i0 += 32 * GF_ITEMS) {
                int64_t j[GF_ITEMS];
                bool have[GF_ITEMS], inr[GF_ITEMS];
#pragma unroll
                for (int q = 0; q < GF_ITEMS; q++) {
                    inr[q] = i0 + q * 32 + lane < len;
                    const IdxT x = inr[q] ? __ldcs(idx + start + i0 + q * 32 + lane) : (IdxT)~(IdxT)0;
                    have[q] = !idx_is_null(x);
                    j[q] = have[q] ? (int64_t)x : 0;
                }
#pragma unroll
                for (int c = 0; c < GF_MAX_COLS; c++) {
                    if (c >= cols.n) break;
                    const GatherCol &gc = cols.c[c];
                    switch (gc.width) {
                        case 1: gather_seg_items<uint8_t>(gc, j, have, inr, start + i0, lane, nulls[c], pol); break;
                        case 2: gather_seg_items<uint16_t>(gc, j, have, inr, start + i0, lane, nulls[c], pol); break;
                        case 4: gather_seg_items<uint32_t>(gc, j, have, inr, start + i0, lane, nulls[c], pol); break;
                        case 8: gather_seg_items<unsigned long long>(gc, j, have, inr, start + i0, lane, nulls[c], pol); break;
                        default: gather_seg_items<ulonglong2>(gc, j, have, inr, start + i0, lane, nulls[c], pol); break;
                    }
                }
            }
        }
    }
#pragma unroll
    for (int c = 0; c < GF_MAX_COLS; c++) {
        if (c >= cols.n) break;
        const unsigned long long z = warp_sum_u32(nulls[c]);
        if (lane == 0 && z) atomicAdd(null_counts + c, z);
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// B200J_OUT_GATHER_HINT: what the dense probe left with the result makes both remaining random reads of the gather go away.
//
// (1) Key columns from the partitioned key stream. The stream the probe walked holds the probe key of every result row IN
//     RESULT ORDER (the result arrays are indexed by stream position), so both key columns of the output are a sequential copy
//     (a widening one when the stream carries 8-byte keys as 32-bit values). Only the build key's validity needs the pairs, and
//     only when some rows have no match. The batch has no NULL keys (the host checks), so the probe key column has no nulls.
// (2) Build payload from slot-ordered copies. slot_order_kernel lays every fixed-width payload column out by TABLE SLOT (entry s
//     = the value of the build row that slot s holds); the probe wrote the slot of every pair next to the build index. A
//     partition's slots are one contiguous range, so its part of a slot-ordered column is a dense array (15 MB per 8-byte
//     column at C2) that stays in L2 while gather_seg_multi_kernel walks the partition's segments -- where the arrival-ordered
//     build store spreads a partition's rows over 780 K different 128 B lines (100 MB of L2 lines for 6 MB of values).
// ---------------------------------------------------------------------------------------------------------------------
template <typename KT, typename T, typename IdxT>
__global__ void __launch_bounds__(256) gather_keys_stream_kernel(const KT *__restrict__ stream, const IdxT *__restrict__ bidx /* null: every row matched */,
                                                                 int64_t n, T *__restrict__ dst_b, uint32_t *__restrict__ dst_b_valid,
                                                                 T *__restrict__ dst_p, unsigned long long *null_b) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    uint32_t nulls = 0;
    for (int64_t row0 = (((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5) * (32 * GF_ITEMS); row0 < n; row0 += warps * (32 * GF_ITEMS)) {
        T v[GF_ITEMS];
        bool ok[GF_ITEMS];
#pragma unroll
        for (int q = 0; q < GF_ITEMS; q++) {
            const int64_t i = row0 + q * 32 + lane;
            v[q] = T{};
            ok[q] = false;
            if (i < n) {
                v[q] = (T)__ldcs(stream + i);
                ok[q] = bidx == nullptr || !idx_is_null(__ldcs(bidx + i));
            }
        }
#pragma unroll
        for (int q = 0; q < GF_ITEMS; q++) {
            const int64_t i = row0 + q * 32 + lane;
            if (i < n) {
                __stcs(dst_p + i, v[q]);
                __stcs(dst_b + i, ok[q] ? v[q] : T{});
                if (!ok[q]) nulls++;
            }
            if (bidx != nullptr) {
                const uint32_t w = __ballot_sync(0xffffffffu, ok[q]);
                if (lane == 0 && row0 + q * 32 < n) dst_b_valid[(row0 + q * 32) >> 5] = w;
            }
        }
    }
    const unsigned long long z = warp_sum_u32(nulls);
    if (lane == 0 && z) atomicAdd(null_b, z);
}

struct SlotOrderCol {
    const void    *src;          // the build store's column (arrival order)
    const uint8_t *src_valid;    // its validity bitmap or null
    void          *dst;          // [n_slots + 2] slot-ordered copy
    uint32_t      *dst_valid;    // its validity words, or null when the column has no nulls
    int32_t        width;
    int32_t        _pad;
};
struct SlotOrderCols {
    SlotOrderCol c[GF_MAX_COLS];
    int32_t      n;
    int32_t      _pad;
};
template <typename T>
__device__ __forceinline__ void slot_order_one(const SlotOrderCol &c, uint64_t s, bool occ, uint32_t row, bool in_range, int lane) {
    bool valid = false;
    if (in_range) {
        T v{};
        if (occ) {
            valid = bit_valid(c.src_valid, (int64_t)row);
            v = reinterpret_cast<const T *>(c.src)[row];
        }
        reinterpret_cast<T *>(c.dst)[s] = v;
    }
    if (c.dst_valid != nullptr) {
        const uint32_t w = __ballot_sync(0xffffffffu, valid);
        if (lane == 0 && in_range) c.dst_valid[s >> 5] = w;
    }
}
// unique build keys only (every occupied slot has count 1 and `row` is the build row)
__global__ void __launch_bounds__(256) slot_order_kernel(const Slot *__restrict__ slots, uint64_t n_slots, const Slot *__restrict__ zero_slot,
                                                         SlotOrderCols cols) {
    const int lane = threadIdx.x & 31;
    const uint64_t total = n_slots + 2, total_round = (total + 31) & ~31ull;
    for (uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; s < total_round; s += (uint64_t)gridDim.x * blockDim.x) {
        const bool in_range = s < total;
        bool occ = false;
        uint32_t row = 0;
        if (in_range) {
            const Slot v = load_slot(s < n_slots ? slots + s : zero_slot + (s - n_slots));
            occ = s < n_slots ? v.key != 0ULL : v.count != 0u;
            row = v.row;
        }
#pragma unroll
        for (int c = 0; c < GF_MAX_COLS; c++) {
            if (c >= cols.n) break;
            switch (cols.c[c].width) {
                case 1: slot_order_one<uint8_t>(cols.c[c], s, occ, row, in_range, lane); break;
                case 2: slot_order_one<uint16_t>(cols.c[c], s, occ, row, in_range, lane); break;
                case 4: slot_order_one<uint32_t>(cols.c[c], s, occ, row, in_range, lane); break;
                case 8: slot_order_one<unsigned long long>(cols.c[c], s, occ, row, in_range, lane); break;
                default: slot_order_one<ulonglong2>(cols.c[c], s, occ, row, in_range, lane); break;
            }
        }
    }
}

__device__ __forceinline__ int64_t utf8_off(const GatherSrc &s, int64_t j) {
    return s.offsets64 ? reinterpret_cast<const int64_t *>(s.offsets)[j] : (int64_t)reinterpret_cast<const int32_t *>(s.offsets)[j];
}

// Utf8 pass 1: byte length of every output row (0 for NULL) + validity word + null count
template <typename IdxT>
__global__ void __launch_bounds__(256) gather_utf8_len_kernel(GatherSrc src, const IdxT *__restrict__ idx, int64_t n, int64_t *__restrict__ len,
                                                              uint32_t *__restrict__ dst_valid, unsigned long long *null_count) {
    const int lane = threadIdx.x & 31;
    uint32_t nulls = 0;
    const int64_t n_round = (n + 31) & ~(int64_t)31;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += (int64_t)gridDim.x * blockDim.x) {
        bool valid = false;
        if (i < n) {
            const IdxT j = idx ? idx[i] : (IdxT)~(IdxT)0;
            int64_t l = 0;
            if (!idx_is_null(j) && src.offsets != nullptr && bit_valid(src.validity, src.bit_offset + (int64_t)j)) {
                l = utf8_off(src, (int64_t)j + 1) - utf8_off(src, (int64_t)j);
                valid = true;
            }
            len[i] = l;
            if (!valid) nulls++;
        }
        const uint32_t word = __ballot_sync(0xffffffffu, valid);
        if (lane == 0 && dst_valid) dst_valid[i >> 5] = word;
    }
    const unsigned long long z = warp_sum_u32(nulls);
    if (lane == 0 && z) atomicAdd(null_count, z);
}

// Exclusive scan of int64 values, three launches: block sums, scan of the block sums (one block), block scan + carry.
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS   = 8;
constexpr int SCAN_TILE    = SCAN_THREADS * SCAN_ITEMS;

__global__ void __launch_bounds__(SCAN_THREADS) scan_block_sums_kernel(const int64_t *__restrict__ in, int64_t n, int64_t *__restrict__ block_sum) {
    __shared__ int64_t s_w[SCAN_THREADS / 32];
    const int64_t base = (int64_t)blockIdx.x * SCAN_TILE;
    int64_t sum = 0;
    for (int q = 0; q < SCAN_ITEMS; q++) {
        const int64_t i = base + (int64_t)q * SCAN_THREADS + threadIdx.x;
        if (i < n) sum += in[i];
    }
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = sum;
    __syncthreads();
    if (threadIdx.x == 0) {
        int64_t t = 0;
        for (int i = 0; i < SCAN_THREADS / 32; i++) t += s_w[i];
        block_sum[blockIdx.x] = t;
    }
}

__global__ void __launch_bounds__(1024) scan_serial_kernel(int64_t *__restrict__ v, int64_t n, int64_t *__restrict__ total) {
    // n = number of tiles (<= a few hundred thousand): chunked serial scan by one block
    __shared__ int64_t s_part[1024];
    const int t = threadIdx.x;
    const int64_t per = (n + 1023) / 1024;
    const int64_t lo = min((int64_t)t * per, n), hi = min(lo + per, n);
    int64_t sum = 0;
    for (int64_t i = lo; i < hi; i++) sum += v[i];
    s_part[t] = sum;
    __syncthreads();
    if (t == 0) {
        int64_t run = 0;
        for (int i = 0; i < 1024; i++) { const int64_t x = s_part[i]; s_part[i] = run; run += x; }
        *total = run;
    }
    __syncthreads();
    int64_t run = s_part[t];
    for (int64_t i = lo; i < hi; i++) { const int64_t x = v[i]; v[i] = run; run += x; }
}

// writes Arrow int32 offsets: out[i] = exclusive prefix, out[n] = total
__global__ void __launch_bounds__(SCAN_THREADS) scan_apply_kernel(const int64_t *__restrict__ in, int64_t n, const int64_t *__restrict__ block_base,
                                                                  int64_t *__restrict__ start64, int32_t *__restrict__ out32) {
    __shared__ int64_t s_w[SCAN_THREADS / 32];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    // thread t owns SCAN_ITEMS consecutive elements
    const int64_t i0 = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)t * SCAN_ITEMS;
    int64_t v[SCAN_ITEMS], sum = 0;
    for (int q = 0; q < SCAN_ITEMS; q++) { v[q] = (i0 + q < n) ? in[i0 + q] : 0; sum += v[q]; }
    int64_t incl = sum;
    for (int o = 1; o < 32; o <<= 1) { const int64_t y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
    if (lane == 31) s_w[warp] = incl;
    __syncthreads();
    int64_t run = block_base[blockIdx.x] + incl - sum;
    for (int w = 0; w < warp; w++) run += s_w[w];
    for (int q = 0; q < SCAN_ITEMS; q++) {
        if (i0 + q < n) { start64[i0 + q] = run; out32[i0 + q] = (int32_t)run; }
        run += v[q];
    }
    if (i0 <= n && n < i0 + SCAN_ITEMS) out32[n] = (int32_t)run;      // the thread whose range holds position n
}

// Utf8 pass 2: a group of LANES lanes per output row copies its bytes (LANES = 32: one warp per row). Short strings (names,
// flags, dates: a dozen bytes) leave most of a warp idle when a whole warp takes one row, so the host picks the group size from
// the average length of the gathered values: 4 lanes up to 6 bytes, 8 up to 16, 16 up to 48, a warp beyond. A group's loads and
// stores are consecutive bytes of one row, i.e. one or two sectors per step either way.
template <typename IdxT, int LANES>
__global__ void __launch_bounds__(256) gather_utf8_copy_kernel(GatherSrc src, const IdxT *__restrict__ idx, int64_t n, const int64_t *__restrict__ start,
                                                               const int64_t *__restrict__ len, char *__restrict__ dst) {
    const int sub = threadIdx.x & (LANES - 1);
    const int64_t groups = ((int64_t)gridDim.x * blockDim.x) / LANES;
    for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / LANES; i < n; i += groups) {
        const int64_t l = len[i];
        if (l == 0) continue;
        const int64_t s = utf8_off(src, (int64_t)idx[i]), d = start[i];
        const char *sp = reinterpret_cast<const char *>(src.values) + s;
        for (int64_t k = sub; k < l; k += LANES) dst[d + k] = sp[k];
    }
}

}  // namespace b200j
