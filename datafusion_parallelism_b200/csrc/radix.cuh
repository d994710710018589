// radix.cuh — shared-memory-staged radix partitioning (histogram + scatter) in front of build and probe.
//
// No counterpart in the reference (its table is one shared concurrent map, version10/new_map_3): this is the
// B200 answer to "the table does not fit L2". Both join sides are split on the TOP radix_bits of the row hash
// (slot selection inside a partition uses the LOW bits, so the two are independent); the partitioned keys are
// processed in partition order, so at any moment the running CTAs touch one or two partitions' tables, which
// stay resident in the 126 MB L2 while the partitioned (key, row id) stream is the only HBM traffic.
//
//   radix_hist_kernel     8 B/tuple read            per-CTA shared-memory histogram -> global histogram
//   radix_scan_kernel     P+1 counters              exclusive scan -> partition offsets, scatter cursors
//   radix_scatter_kernel  8 B read + 12 B write     tile is sorted by partition in shared memory, then each
//                                                   partition's run is written contiguously (coalesced)
// Bin P (one past the last partition) collects NULL-key rows: they never match (inner_join_with_nulls,
// src/lib.rs:149-193) but Right/Full/RightAnti still have to emit them.
#pragma once
#include <type_traits>
#include "kernels.cuh"

namespace b200j {

constexpr int RP_THREADS = 512;
constexpr int RP_ITEMS   = 8;
constexpr int RP_TILE    = RP_THREADS * RP_ITEMS;      // 4096 rows per CTA iteration
constexpr int RP_MAX_BITS = 12;

struct RadixScratch {
    unsigned long long *hist = nullptr;     // [P + 2]
    unsigned long long *off = nullptr;      // [P + 2] exclusive offsets; off[P] = first NULL-key row, off[P+1] = n
    unsigned long long *cursor = nullptr;   // [P + 2]
    unsigned long long *h_off = nullptr;    // pinned host mirror of off[]
    int alloc_bits = -1;
};

struct PartArgs {
    const void         *keys;
    const uint8_t      *validity;
    int64_t             bit_offset;
    int64_t             n;
    int32_t             mis;
    int32_t             radix_bits;
    HashKeys            hk;
    unsigned long long *hist;
    unsigned long long *cursor;
    void               *out_keys;
    uint32_t           *out_rid;
};

__host__ __device__ inline size_t radix_scatter_smem(int kw, int radix_bits) {
    const size_t P1 = ((size_t)1 << radix_bits) + 1;
    return (size_t)RP_TILE * (kw + 4 + 2) + P1 * (4 + 4 + 8) + 64;
}

template <int KW>
__global__ void __launch_bounds__(RP_THREADS, 2) radix_hist_kernel(PartArgs a) {
    extern __shared__ uint32_t s_hist[];
    const int t = threadIdx.x;
    const uint32_t P = 1u << a.radix_bits;
    for (uint32_t p = t; p <= P; p += RP_THREADS) s_hist[p] = 0;
    __syncthreads();
    const int64_t n_tiles = (a.n + a.mis + RP_TILE - 1) / RP_TILE;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        uint64_t key[RP_ITEMS];
        int64_t row[RP_ITEMS];
        load_keys_tile<KW, RP_THREADS, RP_ITEMS>(a.keys, a.mis, a.n, tile, t, key, row);
#pragma unroll
        for (int q = 0; q < RP_ITEMS; q++) {
            if (row[q] < 0 || row[q] >= a.n) continue;
            const uint32_t p = bit_valid(a.validity, a.bit_offset + row[q]) ? (uint32_t)(hash_one(a.hk, key[q]) >> (64 - a.radix_bits)) : P;
            atomicAdd(&s_hist[p], 1u);
        }
    }
    __syncthreads();
    for (uint32_t p = t; p <= P; p += RP_THREADS)
        if (s_hist[p]) atomicAdd(&a.hist[p], (unsigned long long)s_hist[p]);
}

// One block: exclusive scan of hist[0..P] -> off[0..P+1]; cursor[p] = off[p].
__global__ void __launch_bounds__(1024) radix_scan_kernel(const unsigned long long *__restrict__ hist, int n_bins,
                                                          unsigned long long *__restrict__ off, unsigned long long *__restrict__ cursor) {
    __shared__ unsigned long long s_part[1024];
    const int t = threadIdx.x;
    const int per = (n_bins + 1023) / 1024;
    const int lo = t * per, hi = min(lo + per, n_bins);
    unsigned long long sum = 0;
    for (int i = lo; i < hi; i++) sum += hist[i];
    s_part[t] = sum;
    __syncthreads();
    if (t == 0) {
        unsigned long long run = 0;
        for (int i = 0; i < 1024; i++) { const unsigned long long v = s_part[i]; s_part[i] = run; run += v; }
        off[n_bins] = run;
        cursor[n_bins] = run;
    }
    __syncthreads();
    unsigned long long run = s_part[t];
    for (int i = lo; i < hi; i++) {
        off[i] = run;
        cursor[i] = run;
        run += hist[i];
    }
}

template <int KW>
__global__ void __launch_bounds__(RP_THREADS, 2) radix_scatter_kernel(PartArgs a) {
    using KeyT = typename std::conditional<KW == 8, unsigned long long, uint32_t>::type;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const uint32_t P = 1u << a.radix_bits, P1 = P + 1;
    KeyT *s_key = reinterpret_cast<KeyT *>(smem_raw);
    uint32_t *s_rid = reinterpret_cast<uint32_t *>(smem_raw + (size_t)RP_TILE * KW);
    unsigned long long *s_gbase = reinterpret_cast<unsigned long long *>(smem_raw + (size_t)RP_TILE * (KW + 4));
    uint32_t *s_cnt = reinterpret_cast<uint32_t *>(s_gbase + P1);
    uint32_t *s_off = s_cnt + P1;
    uint16_t *s_part = reinterpret_cast<uint16_t *>(s_off + P1);
    __shared__ uint32_t s_wsum[RP_THREADS / 32];
    __shared__ uint32_t s_total;

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int64_t n_tiles = (a.n + a.mis + RP_TILE - 1) / RP_TILE;
    KeyT *out_keys = reinterpret_cast<KeyT *>(a.out_keys);

    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        for (uint32_t p = t; p < P1; p += RP_THREADS) s_cnt[p] = 0;
        __syncthreads();

        uint64_t key[RP_ITEMS];
        int64_t row[RP_ITEMS];
        uint32_t part[RP_ITEMS], rank[RP_ITEMS];
        load_keys_tile<KW, RP_THREADS, RP_ITEMS>(a.keys, a.mis, a.n, tile, t, key, row);
#pragma unroll
        for (int q = 0; q < RP_ITEMS; q++) {
            part[q] = 0xffffffffu;
            if (row[q] < 0 || row[q] >= a.n) continue;
            part[q] = bit_valid(a.validity, a.bit_offset + row[q]) ? (uint32_t)(hash_one(a.hk, key[q]) >> (64 - a.radix_bits)) : P;
            rank[q] = atomicAdd(&s_cnt[part[q]], 1u);
        }
        __syncthreads();

        // exclusive scan of s_cnt[0..P] -> s_off; claim each non-empty partition's run in the global output
        {
            const uint32_t per = (P1 + RP_THREADS - 1) / RP_THREADS;
            const uint32_t lo = t * per, hi = min(lo + per, P1);
            uint32_t sum = 0;
            for (uint32_t i = lo; i < hi; i++) sum += s_cnt[i];
            uint32_t incl = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
            if (lane == 31) s_wsum[warp] = incl;
            __syncthreads();
            if (warp == 0) {
                uint32_t w = lane < RP_THREADS / 32 ? s_wsum[lane] : 0u;
                uint32_t wi = w;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, wi, o); if (lane >= o) wi += y; }
                if (lane < RP_THREADS / 32) s_wsum[lane] = wi - w;
                if (lane == 31) s_total = wi;
            }
            __syncthreads();
            uint32_t run = s_wsum[warp] + incl - sum;
            for (uint32_t i = lo; i < hi; i++) {
                const uint32_t c = s_cnt[i];
                s_off[i] = run;
                if (c) s_gbase[i] = atomicAdd(&a.cursor[i], (unsigned long long)c);
                run += c;
            }
        }
        __syncthreads();

        // sort the tile by partition in shared memory
#pragma unroll
        for (int q = 0; q < RP_ITEMS; q++) {
            if (part[q] == 0xffffffffu) continue;
            const uint32_t pos = s_off[part[q]] + rank[q];
            s_key[pos] = (KeyT)key[q];
            s_rid[pos] = (uint32_t)row[q];
            s_part[pos] = (uint16_t)part[q];
        }
        __syncthreads();

        // write-out: consecutive threads -> consecutive elements of a partition's run -> consecutive addresses
        const uint32_t total = s_total;
        for (uint32_t i = t; i < total; i += RP_THREADS) {
            const uint32_t p = s_part[i];
            const unsigned long long dst = s_gbase[p] + (i - s_off[p]);
            out_keys[dst] = s_key[i];
            a.out_rid[dst] = s_rid[i];
        }
        __syncthreads();
    }
}

// (NULL, probe) rows for the NULL-key probe rows of Right / Full / RightAnti.
__global__ void __launch_bounds__(256) emit_null_rows_kernel(const uint32_t *__restrict__ rid, int64_t n, uint64_t *out_build,
                                                             uint32_t *out_probe, uint64_t at) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        if (out_build) out_build[at + i] = ~0ULL;
        out_probe[at + i] = rid[i];
    }
}

}  // namespace b200j
