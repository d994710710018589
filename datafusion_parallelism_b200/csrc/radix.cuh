// radix.cuh — shared-memory-staged partitioning (histogram + scatter) in front of build and probe, and in
// front of the multi-GPU exchange.
//
// No counterpart in the reference (its table is one shared concurrent map, version10/new_map_3): this is the
// B200 answer to "the table does not fit L2". Both join sides are split on the TOP radix_bits of the row hash
// (slot selection inside a partition uses the LOW 32 bits, the destination GPU of the multi-GPU exchange uses
// bits 32..51, so the three are independent); the partitioned keys are processed in partition order, so at
// any moment the running CTAs touch one or two partitions' tables, which stay resident in the 126 MB L2
// while the partitioned (key, row id) stream is the only HBM traffic.
//
//   part_hist_kernel      8 B/tuple read            CTA b owns a contiguous range of tiles; shared-memory histogram
//                                                   -> cta_hist[bin][b]   (no global atomics anywhere)
//   part_scan_rows_kernel / part_scan_bins_kernel   within-bin exclusive offsets per CTA, bin bases
//   part_scatter_kernel   8 B read + 12 B write     same tile ownership; the tile is sorted by bin in shared memory,
//                                                   then each bin's run is appended to the CTA's own contiguous
//                                                   region of that bin (cursor kept in shared memory)
// Bin n_bins (one past the last) collects NULL-key rows: they never match (inner_join_with_nulls,
// src/lib.rs:149-193) but Right/Full/RightAnti still have to emit them.
//
// DEST mode (multi-GPU): bins are destination ranks, and every bin may have its own output base pointer —
// a peer GPU's receive buffer mapped over NVLink — so the scatter IS the all-to-all (fused partition + exchange).
#pragma once
#include <type_traits>
#include "kernels.cuh"

namespace b200j {

#ifndef B200J_RP_THREADS
#define B200J_RP_THREADS 512
#endif
#ifndef B200J_RP_ITEMS
#define B200J_RP_ITEMS 8
#endif
#ifndef B200J_RP_CTAS
#define B200J_RP_CTAS 2
#endif
constexpr int RP_THREADS = B200J_RP_THREADS;
constexpr int RP_ITEMS   = B200J_RP_ITEMS;
constexpr int RP_CTAS    = B200J_RP_CTAS;            // resident CTAs per SM the kernels are tuned for
constexpr int RP_TILE    = RP_THREADS * RP_ITEMS;      // rows per iteration of the histogram kernel
#ifndef B200J_SC_THREADS
#define B200J_SC_THREADS 512
#endif
#ifndef B200J_SC_TILE
#define B200J_SC_TILE 8192
#endif
#ifndef B200J_SC_CTAS
#define B200J_SC_CTAS 2
#endif
constexpr int SC_THREADS = B200J_SC_THREADS;           // scatter kernel
constexpr int SC_TILE    = B200J_SC_TILE;              // rows per iteration of the scatter kernel; a CTA owns tiles_per_cta of them
#ifndef B200J_SC_NBUF
#define B200J_SC_NBUF 1
#endif
constexpr int SC_NBUF    = B200J_SC_NBUF;              // key staging buffers: 2 = the next tile streams in while this one is processed.
// Measured on B200, 1 B rows, 2^7 bins, histogram + scatter (threads x tile rows x CTAs/SM x staging buffers):
//   512 x 4096 x 2 x 2   7.95 ms   (first version: the next tile streams in while this one is processed)
//   256 x 4096 x 3 x 1   7.55 ms   (a third CTA hides the un-prefetched tile better than a second buffer does)
//   256 x 2048 x 4 x 1  10.0  ms   1024 x 8192 x 1 x 2   8.6 ms   1024 x 16384 x 1 x 1   9.8 ms
//   512 x 8192 x 2 x 1   7.06 ms   and with 512-thread histogram CTAs (2 per SM, same row ranges) 6.78 ms   <- current
// The run a tile appends to a bin (tile / bins = 64 rows = 512 B of keys) must be long for DRAM write locality AND at least
// two CTAs per SM must overlap their barrier phases.
constexpr int SC_CTAS    = B200J_SC_CTAS;              // resident CTAs per SM the scatter kernel (and the grid of both passes) is sized for
static_assert(SC_TILE % RP_TILE == 0, "a scatter tile must be a whole number of histogram tiles");
constexpr int RP_MAX_BITS = 12;
constexpr int RP_MAX_DEST = 16;

struct RadixScratch {
    uint32_t *cta_hist = nullptr;           // [n_bins + 1][grid] counts, scanned in place to within-bin offsets
    uint32_t *bin_tot = nullptr;            // [n_bins + 1]
    uint32_t *bin_base = nullptr;           // [n_bins + 1] base of every bin in the output array (uploaded in DEST mode)
    unsigned long long *off = nullptr;      // [n_bins + 2] exclusive offsets; off[n_bins] = first NULL-key row, off[n_bins+1] = n
    unsigned long long *h_off = nullptr;    // pinned host mirror of off[]
    size_t hist_cap = 0;
    int bins_cap = 0;
};

struct PartArgs {
    const void         *keys;
    const uint8_t      *validity;
    int64_t             bit_offset;
    uint32_t            n;
    uint32_t            mis;
    int32_t             radix_bits;       // radix mode: bin = hash >> (64 - radix_bits)
    int32_t             n_dest;           // DEST mode:  bin = (((hash >> 32) & 0xfffff) * n_dest) >> 20
    uint32_t            n_bins;           // bins without the NULL bin
    uint32_t            tiles_per_cta;    // scatter tiles (SC_TILE rows) owned by one CTA, in both kernels
    HashKeys            hk;
    uint32_t           *cta_hist;
    const uint32_t     *bin_base;
    void               *out_keys;         // radix mode
    uint32_t           *out_rid;
    void               *dest_keys[RP_MAX_DEST + 2];    // DEST mode: output base per bin (may be peer memory)
    uint32_t           *dest_rid[RP_MAX_DEST + 2];
    const unsigned long long *hot_keys;   // DEST mode: open-addressing set of heavy-hitter key values (0 = empty slot) or null
    uint32_t            hot_mask;         // its size - 1
    uint32_t            rid_base;         // added to the local row index (row ids of a chunked side)
    // Single-pass mode (b200j_shuffle_local_scatter): no histogram pass. Every bin owns a region of reserve_cap rows; a tile
    // reserves its run of a bin with ONE atomicAdd on reserve[bin] (a handful of global atomics per 8192 rows). A run that
    // would cross the region's end is dropped and *reserve_ovf set: the caller retries with larger regions.
    uint32_t           *reserve;          // [n_bins + 1] rows reserved so far per bin, or null (two-pass mode: cursors from cta_hist)
    uint32_t            reserve_cap;
    uint32_t           *reserve_ovf;
    // DEST mode, 8-byte keys: the destination buffers take the keys as 32-bit values (half the bytes over NVLink); a key outside
    // [0, 2^32) sets *narrow_ovf and the caller repeats the pass wide
    uint32_t            narrow_out;
    uint32_t           *narrow_ovf;
};


// Heavy-hitter routing of the multi-GPU exchange (SURVEY §8e "skew"): rows whose key is in the hot set are NOT sent to the
// key's owner but kept on this GPU (bin n_dest), where they are joined against the broadcast build rows of the hot keys.
__device__ __forceinline__ bool part_is_hot(const PartArgs &a, uint64_t key) {
    if (key == 0ULL) return false;
    uint32_t i = (uint32_t)((key * 0x9E3779B97F4A7C15ULL) >> 32) & a.hot_mask;
    for (;;) {
        const unsigned long long k = __ldg(a.hot_keys + i);
        if (k == key) return true;
        if (k == 0ULL) return false;
        i = (i + 1u) & a.hot_mask;
    }
}

template <bool DEST>
__device__ __forceinline__ uint32_t part_bin(const PartArgs &a, uint64_t h, uint64_t key) {
    if (DEST) {
        if (a.hot_keys != nullptr && part_is_hot(a, key)) return (uint32_t)a.n_dest;
        return (uint32_t)((((uint32_t)(h >> 32) & 0xfffffu) * (uint32_t)a.n_dest) >> 20);
    }
    return (uint32_t)(h >> (64 - a.radix_bits));
}

// Keys of one partition tile: thread t takes RP_ITEMS * KW / 16 vectors, RP_THREADS vectors apart (a warp's loads
// are 512 contiguous bytes). row = index in the batch; bit q of `in` = item q is a row of the batch.
template <int KW, bool GUARD>
__device__ __forceinline__ void part_load_tile(const PartArgs &a, uint32_t tile, int t, uint64_t (&key)[RP_ITEMS],
                                               uint32_t (&row)[RP_ITEMS], uint32_t &in) {
    constexpr int PER_VEC = 16 / KW;
    constexpr int NVEC = RP_ITEMS / PER_VEC;
    const uint32_t vbase = tile * (uint32_t)RP_TILE;
    const uint32_t mis = a.mis, vend = a.n + mis;
    in = (1u << RP_ITEMS) - 1u;
    if (KW == 8) {
        const unsigned long long *k64 = reinterpret_cast<const unsigned long long *>(a.keys) - mis;
#pragma unroll
        for (int v = 0; v < NVEC; v++) {
            const uint32_t vi = vbase + ((uint32_t)v * RP_THREADS + t) * 2u;
            row[2 * v] = vi - mis;
            row[2 * v + 1] = vi + 1 - mis;
            if (!GUARD || (vi >= mis && vi + 1 < vend)) {
                const ulonglong2 x = __ldcs(reinterpret_cast<const ulonglong2 *>(k64 + vi));
                key[2 * v] = x.x;
                key[2 * v + 1] = x.y;
            } else {
                const bool i0 = vi >= mis && vi < vend, i1 = vi + 1 >= mis && vi + 1 < vend;
                key[2 * v] = i0 ? __ldcs(k64 + vi) : 0ULL;
                key[2 * v + 1] = i1 ? __ldcs(k64 + vi + 1) : 0ULL;
                if (!i0) in &= ~(1u << (2 * v));
                if (!i1) in &= ~(1u << (2 * v + 1));
            }
        }
    } else {
        const uint32_t *k32 = reinterpret_cast<const uint32_t *>(a.keys) - mis;
#pragma unroll
        for (int v = 0; v < NVEC; v++) {
            const uint32_t vi = vbase + ((uint32_t)v * RP_THREADS + t) * 4u;
#pragma unroll
            for (int q = 0; q < 4; q++) row[4 * v + q] = vi + q - mis;
            if (!GUARD || (vi >= mis && vi + 3 < vend)) {
                const uint4 x = __ldcs(reinterpret_cast<const uint4 *>(k32 + vi));
                key[4 * v] = x.x; key[4 * v + 1] = x.y; key[4 * v + 2] = x.z; key[4 * v + 3] = x.w;
            } else {
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    const bool i0 = vi + q >= mis && vi + q < vend;
                    key[4 * v + q] = i0 ? (uint64_t)__ldcs(k32 + vi + q) : 0ULL;
                    if (!i0) in &= ~(1u << (4 * v + q));
                }
            }
        }
    }
}

// bins of one tile; 0xffffffff = item not in the batch
template <int KW, bool DEST, bool GUARD>
__device__ __forceinline__ void part_tile_bins(const PartArgs &a, uint32_t tile, int t, uint64_t (&key)[RP_ITEMS],
                                               uint32_t (&row)[RP_ITEMS], uint32_t (&bin)[RP_ITEMS]) {
    uint32_t in;
    part_load_tile<KW, GUARD>(a, tile, t, key, row, in);
#pragma unroll
    for (int q = 0; q < RP_ITEMS; q++) {
        bin[q] = part_bin<DEST>(a, hash_one(a.hk, key[q]), key[q]);
        if (GUARD) {
            if (!((in >> q) & 1u)) bin[q] = 0xffffffffu;
            else if (a.validity != nullptr && !bit_valid(a.validity, a.bit_offset + (int64_t)row[q])) bin[q] = a.n_bins;
        }
    }
}

__device__ __forceinline__ bool part_tile_is_inner(const PartArgs &a, uint32_t tile) {
    return a.validity == nullptr && (tile > 0 || a.mis == 0) && (uint64_t)(tile + 1) * RP_TILE <= (uint64_t)a.n + a.mis;
}

template <int KW, bool DEST>
__global__ void __launch_bounds__(RP_THREADS, RP_CTAS) part_hist_kernel(PartArgs a) {
    extern __shared__ uint32_t s_hist[];
    const int t = threadIdx.x;
    const uint32_t P1 = a.n_bins + 1;
    for (uint32_t p = t; p < P1; p += RP_THREADS) s_hist[p] = 0;
    __syncthreads();
    const uint32_t n_tiles = (uint32_t)(((uint64_t)a.n + a.mis + RP_TILE - 1) / RP_TILE);
    constexpr uint32_t HT = SC_TILE / RP_TILE;           // histogram tiles per scatter tile: same row range per CTA as the scatter
    const uint32_t t0 = blockIdx.x * a.tiles_per_cta * HT, t1 = min(t0 + a.tiles_per_cta * HT, n_tiles);
    for (uint32_t tile = t0; tile < t1; tile++) {
        uint64_t key[RP_ITEMS];
        uint32_t row[RP_ITEMS], bin[RP_ITEMS];
        if (part_tile_is_inner(a, tile)) {
            part_tile_bins<KW, DEST, false>(a, tile, t, key, row, bin);
#pragma unroll
            for (int q = 0; q < RP_ITEMS; q++) atomicAdd(&s_hist[bin[q]], 1u);
        } else {
            part_tile_bins<KW, DEST, true>(a, tile, t, key, row, bin);
#pragma unroll
            for (int q = 0; q < RP_ITEMS; q++)
                if (bin[q] != 0xffffffffu) atomicAdd(&s_hist[bin[q]], 1u);
        }
    }
    __syncthreads();
    for (uint32_t p = t; p < P1; p += RP_THREADS) a.cta_hist[(size_t)p * gridDim.x + blockIdx.x] = s_hist[p];
}

// One CTA per bin: exclusive scan of the bin's row of per-CTA counts (in place) -> within-bin offsets; bin total.
__global__ void __launch_bounds__(256) part_scan_rows_kernel(uint32_t *__restrict__ cta_hist, uint32_t grid, uint32_t *__restrict__ bin_tot) {
    __shared__ uint32_t s_w[8];
    __shared__ uint32_t s_carry;
    uint32_t *rowp = cta_hist + (size_t)blockIdx.x * grid;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    if (t == 0) s_carry = 0;
    __syncthreads();
    for (uint32_t i0 = 0; i0 < grid; i0 += 256) {
        const uint32_t i = i0 + t;
        const uint32_t c = i < grid ? rowp[i] : 0u;
        uint32_t incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
        if (lane == 31) s_w[warp] = incl;
        __syncthreads();
        uint32_t pre = s_carry;
        for (int w = 0; w < warp; w++) pre += s_w[w];
        if (i < grid) rowp[i] = pre + incl - c;
        __syncthreads();
        if (t == 255) s_carry = pre + incl;
        __syncthreads();
    }
    if (t == 0) bin_tot[blockIdx.x] = s_carry;
}

// One block: exclusive scan of bin_tot[0..n_bins1) -> off[0..n_bins1] (u64, for the host), bin_base (u32, may be null).
__global__ void __launch_bounds__(1024) part_scan_bins_kernel(const uint32_t *__restrict__ bin_tot, int n_bins1,
                                                              unsigned long long *__restrict__ off, uint32_t *__restrict__ bin_base) {
    __shared__ unsigned long long s_part[1024];
    const int t = threadIdx.x;
    const int per = (n_bins1 + 1023) / 1024;
    const int lo = min(t * per, n_bins1), hi = min(lo + per, n_bins1);
    unsigned long long sum = 0;
    for (int i = lo; i < hi; i++) sum += bin_tot[i];
    s_part[t] = sum;
    __syncthreads();
    if (t == 0) {
        unsigned long long run = 0;
        for (int i = 0; i < 1024; i++) { const unsigned long long v = s_part[i]; s_part[i] = run; run += v; }
        off[n_bins1] = run;
    }
    __syncthreads();
    unsigned long long run = s_part[t];
    for (int i = lo; i < hi; i++) {
        off[i] = run;
        if (bin_base) bin_base[i] = (uint32_t)run;
        run += bin_tot[i];
    }
}

// ---- bulk-async (TMA) helpers: one elected thread streams a whole key tile into shared memory, completion on an mbarrier
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_load(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src),
                 "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok)
                     : "r"(smem_u32(bar)), "r"(parity)
                     : "memory");
    } while (!ok);
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }


__host__ __device__ inline size_t part_scatter_smem(int kw, uint32_t n_bins) {
    const size_t P1 = (size_t)n_bins + 1;
    return (size_t)SC_NBUF * SC_TILE * kw + (size_t)SC_TILE * 4 + P1 * 16 + 128;
}

// Scatter of one CTA's row range. Key tiles are double-buffered in shared memory by bulk-async copies (the next tile
// streams in while this one is processed, so no warp ever waits on DRAM and no key lives in a register across a
// barrier); a tile is ranked with shared-memory atomics, its permutation (bin, source index) is laid out bin by bin,
// and the write-out walks that permutation: consecutive threads -> consecutive elements of a bin's run -> consecutive
// global addresses, appended at the CTA's own cursor of that bin (cursors live in shared memory: no global atomics).
template <int KW, bool DEST>
__global__ void __launch_bounds__(SC_THREADS, SC_CTAS) part_scatter_kernel(PartArgs a) {
    using KeyT = typename std::conditional<KW == 8, unsigned long long, uint32_t>::type;
    constexpr int PER_VEC = 16 / KW;
    constexpr int ITEMS = SC_TILE / SC_THREADS;
    constexpr int NVEC = ITEMS / PER_VEC;
    constexpr uint32_t TILE_BYTES = (uint32_t)SC_TILE * KW;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const uint32_t P1 = a.n_bins + 1;
    KeyT *s_stage0 = reinterpret_cast<KeyT *>(smem_raw);
    KeyT *s_stage1 = reinterpret_cast<KeyT *>(smem_raw + (SC_NBUF - 1) * TILE_BYTES);
    uint32_t *s_perm = reinterpret_cast<uint32_t *>(smem_raw + SC_NBUF * TILE_BYTES);     // (bin << 16 | source index), bin by bin
    uint32_t *s_cnt = s_perm + SC_TILE;
    uint32_t *s_off = s_cnt + P1;        // position of the bin's run inside the permuted tile
    uint32_t *s_cur = s_off + P1;        // this CTA's append cursor inside the bin's output
    uint32_t *s_delta = s_cur + P1;      // output position of permuted element i of bin p = s_delta[p] + i
    __shared__ __align__(8) uint64_t s_bar[2];
    __shared__ uint32_t s_wsum[SC_THREADS / 32];
    __shared__ uint32_t s_total;
    __shared__ uint32_t s_drop;          // bit p: this tile's run of bin p did not fit its region (single-pass mode)
    __shared__ uint32_t s_wtot[SC_THREADS / 32][12];   // packed DEST ranking: rows per (warp, bin), then the warp's base inside the bin
    __shared__ KeyT *s_dkeys[RP_MAX_DEST + 2];
    __shared__ uint32_t *s_drid[RP_MAX_DEST + 2];

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const uint32_t mis = a.mis;
    const uint32_t vend = a.n + mis;                                   // end of the aligned "virtual" row space
    const uint32_t n_tiles = (uint32_t)(((uint64_t)vend + SC_TILE - 1) / SC_TILE);
    const uint32_t t0 = blockIdx.x * a.tiles_per_cta, t1 = min(t0 + a.tiles_per_cta, n_tiles);
    const bool reserve = a.reserve != nullptr;
    for (uint32_t p = t; p < P1; p += SC_THREADS) {
        s_cnt[p] = 0;
        s_cur[p] = reserve ? 0u : a.bin_base[p] + a.cta_hist[(size_t)p * gridDim.x + blockIdx.x];
    }
    if (t == 0) s_drop = 0;
    if (DEST && t <= RP_MAX_DEST + 1) { s_dkeys[t] = reinterpret_cast<KeyT *>(a.dest_keys[t]); s_drid[t] = a.dest_rid[t]; }
    if (t == 0) { mbar_init(&s_bar[0], 1); mbar_init(&s_bar[1], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();
    if (t0 >= t1) return;
    const KeyT *gkeys = reinterpret_cast<const KeyT *>(a.keys) - mis;   // virtual row v lives at gkeys[v]; 16 B aligned
    KeyT *out_keys = reinterpret_cast<KeyT *>(a.out_keys);

    // a tile is "inner" when it lies fully inside the batch and there is no validity bitmap: then it is one bulk copy
    auto is_inner = [&](uint32_t tile) {
        return a.validity == nullptr && (tile > 0 || mis == 0) && (uint64_t)(tile + 1) * SC_TILE <= (uint64_t)vend;
    };
    // fills a staging buffer: bulk-async for inner tiles (thread 0 issues, completion on the mbarrier), guarded loads otherwise
    auto fetch = [&](uint32_t tile, int buf) {
        KeyT *dst = buf ? s_stage1 : s_stage0;
        if (is_inner(tile)) {
            if (t == 0) {
                fence_proxy_async();
                mbar_expect_tx(&s_bar[buf], TILE_BYTES);
                bulk_load(dst, gkeys + (size_t)tile * SC_TILE, TILE_BYTES, &s_bar[buf]);
            }
        } else {
            for (uint32_t i = t; i < (uint32_t)SC_TILE; i += SC_THREADS) {
                const uint32_t v = tile * (uint32_t)SC_TILE + i;
                dst[i] = (v >= mis && v < vend) ? __ldcs(gkeys + v) : (KeyT)0;
            }
        }
    };
    bool wide_seen = false;
    const bool pk_mode = DEST && P1 <= 10u && ITEMS <= 63;
    const bool narrow_out = DEST && KW == 8 && a.narrow_out != 0u;
    uint32_t n_bulk[2] = {0, 0};         // bulk copies issued so far per buffer (mbarrier phase parity)
    fetch(t0, 0);
    for (uint32_t tile = t0; tile < t1; tile++) {
        const int buf = SC_NBUF == 2 ? (int)((tile - t0) & 1u) : 0;
        const KeyT *stage = buf ? s_stage1 : s_stage0;
        const bool inner = is_inner(tile);
        // the other buffer was last read in the write-out of tile - 1, which ended with a block barrier: refill it now
        if (SC_NBUF == 2 && tile + 1 < t1) fetch(tile + 1, buf ^ 1);
        if (inner) { mbar_wait(&s_bar[buf], n_bulk[buf] & 1u); n_bulk[buf]++; }
        else __syncthreads();
        // note: n_bulk of the buffer being prefetched is advanced when that tile is consumed

        // ---- rank: bin of every key, rank inside the bin from a shared-memory atomic
        uint32_t packed[ITEMS];
        unsigned long long pcnt = 0;          // DEST mode, <= 10 bins: this thread's rows per bin so far, 6 bits each (a thread has 16 rows)
        const uint32_t vbase = tile * (uint32_t)SC_TILE;
#pragma unroll
        for (int v = 0; v < NVEC; v++) {
            const uint32_t src0 = ((uint32_t)v * SC_THREADS + t) * PER_VEC;
            KeyT k[PER_VEC];
            if (KW == 8) {
                const ulonglong2 x = *reinterpret_cast<const ulonglong2 *>(stage + src0);
                k[0] = x.x; k[1] = x.y;
            } else {
                const uint4 x = *reinterpret_cast<const uint4 *>(stage + src0);
                k[0] = x.x; k[1] = x.y; k[2] = x.z; k[3] = x.w;
            }
#pragma unroll
            for (int j = 0; j < PER_VEC; j++) {
                uint32_t bin = part_bin<DEST>(a, hash_one(a.hk, (uint64_t)k[j]), (uint64_t)k[j]);
                if (DEST && KW == 8) wide_seen |= ((unsigned long long)k[j] >> 32) != 0ULL;
                bool in = true;
                if (!inner) {
                    const uint32_t vrow = vbase + src0 + j;
                    in = vrow >= mis && vrow < vend;
                    if (in && a.validity != nullptr && !bit_valid(a.validity, a.bit_offset + (int64_t)(vrow - mis))) bin = a.n_bins;
                }
                if (DEST && pk_mode) {
                    // up to 10 bins (8 ranks + NULL + hot): no shared memory, no vote at all while ranking. Every thread counts its
                    // own rows per bin in ONE 64-bit register (6-bit fields); the per-bin prefix over the threads follows below.
                    const uint32_t sh = bin * 6u;
                    const uint32_t rk = (uint32_t)(pcnt >> sh) & 63u;
                    if (in) pcnt += 1ULL << sh;
                    packed[v * PER_VEC + j] = in ? ((bin << 16) | rk) : 0xffffffffu;
                } else if (DEST) {
                    // few bins (ranks + NULL + hot <= 18): a per-lane shared-memory atomic would serialise 32 lanes on a handful of
                    // addresses (2 GPUs: 16-way). The lanes of a bin are matched with 5 ballots and ONE lane adds their number.
                    const uint32_t mb = in ? bin : 31u;
                    uint32_t peers = 0xffffffffu;
#pragma unroll
                    for (int bit = 0; bit < 5; bit++) {
                        const bool one = (mb >> bit) & 1u;
                        const uint32_t bl = __ballot_sync(0xffffffffu, one);
                        peers &= one ? bl : ~bl;
                    }
                    const int leader = __ffs(peers) - 1;
                    uint32_t base = 0;
                    if (lane == leader && in) base = atomicAdd(&s_cnt[bin], (uint32_t)__popc(peers));
                    base = __shfl_sync(0xffffffffu, base, leader);
                    packed[v * PER_VEC + j] = in ? ((bin << 16) | (base + __popc(peers & ((1u << lane) - 1u)))) : 0xffffffffu;
                } else {
                    packed[v * PER_VEC + j] = in ? ((bin << 16) | atomicAdd(&s_cnt[bin], 1u)) : 0xffffffffu;
                }
            }
        }
        // packed mode: exclusive prefix of the per-thread counts inside the warp (12-bit fields: bins 0..4 in pre_lo, 5..9 in
        // pre_hi), warp totals to shared memory, then one short pass turns them into per-warp bases and the tile's histogram
        unsigned long long pre_lo = 0, pre_hi = 0;
        if (DEST && pk_mode) {
            for (uint32_t b = 0; b < P1; b++) {
                const uint32_t c = (uint32_t)(pcnt >> (6u * b)) & 63u;
                uint32_t incl = c;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
                if (lane == 31) s_wtot[warp][b] = incl;
                const unsigned long long e = (unsigned long long)(incl - c);
                if (b < 5u) pre_lo |= e << (12u * b); else pre_hi |= e << (12u * (b - 5u));
            }
        }
        __syncthreads();
        if (DEST && pk_mode) {
            if ((uint32_t)t < P1) {
                uint32_t run = 0;
#pragma unroll
                for (int w = 0; w < SC_THREADS / 32; w++) { const uint32_t c = s_wtot[w][t]; s_wtot[w][t] = run; run += c; }
                s_cnt[t] = run;
            }
            __syncthreads();
        }

        // ---- exclusive scan of s_cnt[0..P1) -> s_off; advance the CTA's cursors; clear the counts for the next tile
        {
            const uint32_t per = (P1 + SC_THREADS - 1) / SC_THREADS;
            const uint32_t lo = min(t * per, P1), hi = min(lo + per, P1);
            uint32_t sum = 0;
            for (uint32_t i = lo; i < hi; i++) sum += s_cnt[i];
            uint32_t incl = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
            if (lane == 31) s_wsum[warp] = incl;
            __syncthreads();
            if (warp == 0) {
                uint32_t w = lane < SC_THREADS / 32 ? s_wsum[lane] : 0u;
                uint32_t wi = w;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, wi, o); if (lane >= o) wi += y; }
                if (lane < SC_THREADS / 32) s_wsum[lane] = wi - w;
                if (lane == 31) s_total = wi;
            }
            __syncthreads();
            uint32_t run = s_wsum[warp] + incl - sum;
            for (uint32_t i = lo; i < hi; i++) {
                const uint32_t c = s_cnt[i];
                uint32_t cur = s_cur[i];
                if (reserve) {
                    cur = c ? atomicAdd(a.reserve + i, c) : 0u;
                    if (c && (cur > a.reserve_cap || c > a.reserve_cap - cur)) { atomicOr(&s_drop, 1u << (i & 31u)); atomicExch(a.reserve_ovf, 1u); }
                }
                s_off[i] = run;
                s_delta[i] = cur - run;
                s_cur[i] = cur + c;
                s_cnt[i] = 0;
                run += c;
            }
        }
        __syncthreads();

        // ---- permutation: element (bin, rank) sits at s_off[bin] + rank and remembers where its key is staged
#pragma unroll
        for (int q = 0; q < ITEMS; q++) {
            if (!inner && packed[q] == 0xffffffffu) continue;
            const uint32_t bin = packed[q] >> 16;
            uint32_t rank = packed[q] & 0xffffu;
            if (DEST && pk_mode) {      // + the rows of this bin in the warps and in the lanes in front of this thread
                const unsigned long long pw = bin < 5u ? pre_lo : pre_hi;
                rank += s_wtot[warp][bin] + ((uint32_t)(pw >> (12u * (bin < 5u ? bin : bin - 5u))) & 4095u);
            }
            const uint32_t src = ((uint32_t)(q / PER_VEC) * SC_THREADS + t) * PER_VEC + (q % PER_VEC);
            s_perm[s_off[bin] + rank] = (bin << 16) | src;
        }
        __syncthreads();

        // ---- write-out
        const uint32_t total = inner ? (uint32_t)SC_TILE : s_total;
        const uint32_t rid0 = vbase - mis + a.rid_base;
        const uint32_t drop = s_drop;
        // unrolled by the whole tile (16 rows per thread): the shared-memory reads of all iterations are in flight before the
        // first global store. 1 B rows, 2^7 bins, histogram + scatter: unroll 2 6.78, 4 6.82, 8 6.40, 16 6.35 ms (63 -> 64
        // registers, no spills); an L2 prefetch of the next tile on top of it bought nothing (profiles/r1_l_scatter_unroll.txt)
#pragma unroll 16
        for (uint32_t i = t; i < total; i += SC_THREADS) {
            const uint32_t pm = s_perm[i];
            const uint32_t p = pm >> 16, src = pm & 0xffffu;
            const uint32_t dst = s_delta[p] + i;
            if (DEST) {
                if (drop && ((drop >> (p & 31u)) & 1u)) continue;
                if (narrow_out) reinterpret_cast<uint32_t *>(s_dkeys[p])[dst] = (uint32_t)stage[src];
                else s_dkeys[p][dst] = stage[src];
                s_drid[p][dst] = rid0 + src;
            } else {
                out_keys[dst] = stage[src];
                a.out_rid[dst] = rid0 + src;
            }
        }
        __syncthreads();
        if (reserve && t == 0) s_drop = 0;
        if (SC_NBUF == 1 && tile + 1 < t1) fetch(tile + 1, 0);
    }
    if (narrow_out && __any_sync(0xffffffffu, wide_seen) && lane == 0) atomicExch(a.narrow_ovf, 1u);
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
