// radix_block.cuh — single-read, block-local radix partitioning of a probe batch (round 2).
//
// The two-pass partition of radix.cuh reads the key stream twice (histogram, then scatter: 28 B per row) because a row's
// position in a globally partitioned array depends on the histogram of ALL rows. The dense probe does not need a globally
// partitioned array: it needs (a) every row at an exact position of a DENSE array, because the row-id array IS the
// result's probe-index column, and (b) the rows of one partition reachable together, so that the partition's table stays
// in L2 while they are probed. Both hold when every BLOCK of PB_CLUSTER x 8192 consecutive rows is partitioned in place:
//
//   block b of the batch  ->  rows [b*B, (b+1)*B) of out_keys / out_rid, sorted by partition (NULL keys last)
//   seg[b][p]             ->  where partition p starts inside block b       (u32 table, [n_blocks][P + 2])
//
// and the probe walks, for p = 0 .. P-1, the segments (p, b) of all blocks (probe_dense_seg_kernel, probe.cuh).
// One read of the keys (8 B) + one write of key and row id (12 B) per row: 20 B instead of 28 B, no histogram kernel,
// no scan kernels, no global cursors.
//
// A block is owned by a thread-block CLUSTER of PB_CLUSTER CTAs (sm_90+: CTAs co-scheduled on one GPC with distributed
// shared memory). Every CTA ranks its own PB_TILE-row tile in shared memory; the CTAs exchange their tile histograms through
// DSMEM (one cluster barrier per tile) and each then knows where its runs go inside the block. Segments are therefore
// PB_BLOCK / 2^radix_bits rows long on average (8 x 4096 rows, 2^7 partitions: 256 rows = 2 KB of keys), long enough for
// the probe to stream them.
//
// Measured on B200, 1 B int64 rows, 2^7 partitions (profiles/r2_b_block_partition.txt), partition pass / segment probe:
//   two-pass global partition + probe_dense_kernel (round 1)      6.37 / 6.59 ms
//   512 threads x 8192-row tiles x 2 CTAs/SM, cluster 4           5.51 / 7.18 ms
//   256 threads x 4096-row tiles x 4 CTAs/SM, cluster 8           5.28 / 7.19 ms   <- default
//   same, cluster 16 (64 K-row blocks, 512-row segments)          6.26 / 6.89 ms
//   128 threads x 2048-row tiles x 8 CTAs/SM, cluster 16          7.11 / 7.19 ms
// Ranking inside a tile: shared-memory atomics on CTA-wide counters (B200J_PB_RANK 2). Warp-synchronous ranking with
// warp-private counters was measured and is SLOWER here: ballot match over the bin's bits (CUB onesweep's WarpMatch) 9.3 ms,
// MATCH.ANY 9.7 ms for the same pass -- the match costs ~70 instructions per row against one ATOMS.
#pragma once
#include "radix.cuh"

namespace b200j {

#ifndef B200J_PB_CLUSTER
#define B200J_PB_CLUSTER 8
#endif
constexpr int PB_CLUSTER = B200J_PB_CLUSTER;          // CTAs per cluster = tiles per block
#ifndef B200J_PB_THREADS
#define B200J_PB_THREADS 256
#endif
#ifndef B200J_PB_TILE
#define B200J_PB_TILE 4096
#endif
#ifndef B200J_PB_CTAS
#define B200J_PB_CTAS 4
#endif
constexpr int PB_THREADS = B200J_PB_THREADS;
constexpr int PB_WARPS   = PB_THREADS / 32;
constexpr int PB_TILE    = B200J_PB_TILE;              // rows per CTA tile
constexpr int PB_CTAS    = B200J_PB_CTAS;              // resident CTAs per SM the kernel is sized for
constexpr int PB_BLOCK   = PB_TILE * PB_CLUSTER;       // rows per block
constexpr int PB_ITEMS   = PB_TILE / PB_THREADS;       // 16
constexpr int PB_MAX_BITS = 8;
static_assert(PB_TILE <= 65536, "permutation entries hold the source index in 16 bits");

struct BlockPartArgs {
    const void         *keys;
    const uint8_t      *validity;
    int64_t             bit_offset;
    uint32_t            n;
    uint32_t            mis;              // keys misaligned from 16 B by this many elements ("virtual" row v = row + mis)
    int32_t             radix_bits;
    uint32_t            n_blocks;
    HashKeys            hk;
    void               *out_keys;
    uint32_t           *out_rid;
    uint32_t           *seg;              // [n_blocks][P + 2]: starts of bins 0 .. P (P = NULL keys) inside the block, then the block's row count
    uint32_t           *block_base;       // [n_blocks] first output row of every block (written here, read by the probe)
    unsigned long long *part_rows;        // [P + 1] rows per partition over the whole batch (atomics, one per bin per block) or null
    uint32_t            rid_base;
    uint32_t           *narrow_ovf;       // NARROW: set to 1 when a key does not fit 32 bits (the caller then repeats the pass wide)
    uint32_t            out_base;         // rows of the pieces in front of this one (a batch may arrive as several pieces, e.g. one per source
                                          // rank of the multi-GPU exchange): this piece's blocks start there in out_keys / out_rid
};

// first output row of block b: blocks are cut in the 16 B-aligned virtual row space, so block 0 is short by `mis` rows
__host__ __device__ __forceinline__ uint32_t pb_block_base(uint32_t b, uint32_t mis) { return b == 0 ? 0u : b * (uint32_t)PB_BLOCK - mis; }

__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ uint32_t cluster_id_x() { uint32_t r; asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r)); return r; }
__device__ __forceinline__ uint32_t cluster_count_x() { uint32_t r; asm volatile("mov.u32 %0, %%nclusterid.x;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
__device__ __forceinline__ uint32_t dsmem_ld_u32(const uint32_t *local, uint32_t rank) {
    uint32_t ra, v;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(smem_u32(local)), "r"(rank));
    asm volatile("ld.shared::cluster.u32 %0, [%1];" : "=r"(v) : "r"(ra) : "memory");
    return v;
}

// lanes of the warp whose `bin` equals mine (all 32 lanes take part).
//   B200J_PB_RANK 0  ballot match over the low `nbits` bits, unrolled for the common bit counts (CUB onesweep's WarpMatch)
//   B200J_PB_RANK 1  the MATCH.ANY instruction
//   B200J_PB_RANK 2  no match at all: shared-memory atomics on CTA-wide counters, as part_scatter_kernel does
#ifndef B200J_PB_RANK
#define B200J_PB_RANK 2
#endif
template <int NBITS>
__device__ __forceinline__ uint32_t match_bits(uint32_t bin) {
    uint32_t peers = 0xffffffffu;
#pragma unroll
    for (int i = 0; i < NBITS; i++) {
        const bool bit = (bin >> i) & 1u;
        const uint32_t b = __ballot_sync(0xffffffffu, bit);
        peers &= bit ? b : ~b;
    }
    return peers;
}
__device__ __forceinline__ uint32_t match_bin(uint32_t bin, int nbits) {
#if B200J_PB_RANK == 1
    (void)nbits;
    return __match_any_sync(0xffffffffu, bin);
#else
    switch (nbits) {
        case 7: return match_bits<7>(bin);
        case 8: return match_bits<8>(bin);
        case 9: return match_bits<9>(bin);
        case 6: return match_bits<6>(bin);
        case 5: return match_bits<5>(bin);
        default: break;
    }
    uint32_t peers = 0xffffffffu;
    for (int i = 0; i < nbits; i++) {
        const bool bit = (bin >> i) & 1u;
        const uint32_t b = __ballot_sync(0xffffffffu, bit);
        peers &= bit ? b : ~b;
    }
    return peers;
#endif
}

__host__ __device__ inline size_t part_block_smem(int kw, uint32_t n_bins) {
    const size_t Q = (size_t)n_bins + 2;                 // bins + NULL bin + the "not a row" bin of ragged tiles
    return (size_t)PB_TILE * kw + (size_t)PB_TILE * 4 + (size_t)PB_WARPS * Q * 2 + Q * 4 * 5 + 256;
}

// NARROW (KW == 8 only): the partitioned stream carries the keys as 32-bit values. Join keys are mostly small non-negative
// integers (ids, surrogate keys): 4 B instead of 8 B per row written here and read again by the probe, which zero-extends
// them before hashing and comparing, so nothing changes for the table or the result. A key outside [0, 2^32) raises
// *narrow_ovf; the host then repeats the pass wide and stops trying for this join (b200join.cu: narrow_ok).
template <int KW, bool NARROW = false>
__global__ void __launch_bounds__(PB_THREADS, PB_CTAS) part_block_kernel(BlockPartArgs a) {
    static_assert(!NARROW || KW == 8, "only 8-byte keys are narrowed");
    using KeyT = typename std::conditional<KW == 8, unsigned long long, uint32_t>::type;
    using OutT = typename std::conditional<(KW == 8 && !NARROW), unsigned long long, uint32_t>::type;
    constexpr int PER_VEC = 16 / KW;
    constexpr int NVEC = PB_ITEMS / PER_VEC;
    constexpr uint32_t TILE_BYTES = (uint32_t)PB_TILE * KW;
    constexpr uint32_t FULL = 0xffffffffu;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const uint32_t P = 1u << a.radix_bits;
    const uint32_t P1 = P + 1;                           // + NULL bin
    const uint32_t Q = P + 2;                            // + "not a row of the batch" (ragged first / last tile)
    KeyT *s_stage = reinterpret_cast<KeyT *>(smem_raw);
    uint32_t *s_perm = reinterpret_cast<uint32_t *>(smem_raw + TILE_BYTES);       // (bin << 16 | source index), bin by bin
    unsigned short *s_wcnt = reinterpret_cast<unsigned short *>(s_perm + PB_TILE);  // [PB_WARPS][Q] rows of (warp, bin), then the warp's offset inside the bin
    uint32_t *s_hist = reinterpret_cast<uint32_t *>(s_wcnt + (size_t)PB_WARPS * Q + ((PB_WARPS * Q) & 1u));   // [2][Q] this tile's histogram (double-buffered: peers read it)
    uint32_t *s_off = s_hist + 2 * Q;                    // [Q] start of the bin's run inside the sorted tile
    uint32_t *s_delta = s_off + Q;                       // [Q] output position of sorted element i of bin p = s_delta[p] + i
    uint32_t *s_tot = s_delta + Q;                       // [Q] rows of the bin in the whole block
    __shared__ __align__(8) uint64_t s_bar;
    __shared__ uint32_t s_scan[64];

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const uint32_t mis = a.mis, vend = a.n + mis;
    const uint32_t n_tiles = (uint32_t)(((uint64_t)vend + PB_TILE - 1) / PB_TILE);
    const uint32_t crank = cluster_ctarank();
    const KeyT *gkeys = reinterpret_cast<const KeyT *>(a.keys) - mis;
    OutT *out_keys = reinterpret_cast<OutT *>(a.out_keys);
    bool wide_seen = false;
    unsigned short *my_wcnt = s_wcnt + (size_t)warp * Q;

    if (t == 0) mbar_init(&s_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();

    auto is_inner = [&](uint32_t tile) {
        return a.validity == nullptr && (tile > 0 || mis == 0) && (uint64_t)(tile + 1) * PB_TILE <= (uint64_t)vend;
    };
    auto fetch = [&](uint32_t tile) {
        if (tile >= n_tiles) return;
        if (is_inner(tile)) {
            if (t == 0) {
                fence_proxy_async();
                mbar_expect_tx(&s_bar, TILE_BYTES);
                bulk_load(s_stage, gkeys + (size_t)tile * PB_TILE, TILE_BYTES, &s_bar);
            }
        } else {
            for (uint32_t i = t; i < (uint32_t)PB_TILE; i += PB_THREADS) {
                const uint32_t v = tile * (uint32_t)PB_TILE + i;
                s_stage[i] = (v >= mis && v < vend) ? __ldcs(gkeys + v) : (KeyT)0;
            }
        }
    };
    uint32_t n_bulk = 0, it = 0;
    const uint32_t blk0 = cluster_id_x(), blk_step = cluster_count_x();
    fetch(blk0 * PB_CLUSTER + crank);
    for (uint32_t blk = blk0; blk < a.n_blocks; blk += blk_step, it++) {
        const uint32_t tile = blk * PB_CLUSTER + crank;
        const bool have = tile < n_tiles;
        const bool inner = have && is_inner(tile);
        const int nbits = inner ? a.radix_bits : a.radix_bits + 1;        // ragged / nullable tiles use bins P and P + 1 too
        uint32_t *hist = s_hist + (it & 1u) * Q;
#if B200J_PB_RANK == 2
        for (uint32_t i = t; i < Q; i += PB_THREADS) s_tot[i] = 0;
#else
        for (uint32_t i = t; i < PB_WARPS * Q; i += PB_THREADS) s_wcnt[i] = 0;
#endif
        if (inner) { mbar_wait(&s_bar, n_bulk & 1u); n_bulk++; }
        __syncthreads();

        // ---- rank: bin of every key; rank inside (warp, bin) from a ballot match, the warp's counters are private to it
        uint32_t packed[PB_ITEMS];
        const uint32_t vbase = tile * (uint32_t)PB_TILE;
        if (have) {
#pragma unroll
            for (int v = 0; v < NVEC; v++) {
                const uint32_t src0 = ((uint32_t)v * PB_THREADS + t) * PER_VEC;
                KeyT k[PER_VEC];
                if (KW == 8) {
                    const ulonglong2 x = *reinterpret_cast<const ulonglong2 *>(s_stage + src0);
                    k[0] = x.x; k[1] = x.y;
                } else {
                    const uint4 x = *reinterpret_cast<const uint4 *>(s_stage + src0);
                    k[0] = x.x; k[1] = x.y; k[2] = x.z; k[3] = x.w;
                }
#pragma unroll
                for (int j = 0; j < PER_VEC; j++) {
                    if (NARROW) wide_seen |= ((unsigned long long)k[j] >> 32) != 0ULL;
                    uint32_t bin = (uint32_t)(hash_one(a.hk, (uint64_t)k[j]) >> (64 - a.radix_bits));
                    if (!inner) {
                        const uint32_t vrow = vbase + src0 + j;
                        if (vrow < mis || vrow >= vend) bin = P1;
                        else if (a.validity != nullptr && !bit_valid(a.validity, a.bit_offset + (int64_t)(vrow - mis))) bin = P;
                    }
#if B200J_PB_RANK == 2
                    packed[v * PER_VEC + j] = (bin << 16) | atomicAdd(&s_tot[bin], 1u);     // s_tot doubles as the CTA-wide counters
#else
                    const uint32_t peers = match_bin(bin, nbits);
                    const uint32_t before = my_wcnt[bin];                  // every peer reads the same counter ...
                    __syncwarp();
                    if ((peers & lt_mask) == 0u) my_wcnt[bin] = (unsigned short)(before + __popc(peers));   // ... its lowest lane advances it
                    __syncwarp();
                    packed[v * PER_VEC + j] = (bin << 16) | (before + __popc(peers & lt_mask));
#endif
                }
            }
        }
        __syncthreads();

        // ---- per bin: exclusive prefix over the warps (-> the warp's offset inside the bin's run), the tile's histogram
        for (uint32_t b = t; b < Q; b += PB_THREADS) {
#if B200J_PB_RANK == 2
            const uint32_t run = s_tot[b];
#else
            uint32_t run = 0;
#pragma unroll
            for (int w = 0; w < PB_WARPS; w++) {
                const uint32_t c = s_wcnt[(size_t)w * Q + b];
                s_wcnt[(size_t)w * Q + b] = (unsigned short)run;
                run += c;
            }
#endif
            hist[b] = b < P1 ? run : 0u;                       // "not a row" entries are dropped here
        }
        __syncthreads();
        cluster_arrive();                                       // my histogram is published

        // ---- exclusive scan of the tile's histogram over the bins (warp 0) while the peers arrive
        if (warp == 0) {
            const uint32_t per = (P1 + 31) / 32;
            const uint32_t lo = min((uint32_t)lane * per, P1), hi = min(lo + per, P1);
            uint32_t sum = 0;
            for (uint32_t i = lo; i < hi; i++) sum += hist[i];
            uint32_t incl = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += y; }
            uint32_t run = incl - sum;
            for (uint32_t i = lo; i < hi; i++) { s_off[i] = run; run += hist[i]; }
            if (lane == 31) s_scan[0] = incl;                  // rows of this tile
        }
        cluster_wait();                                         // every CTA of the cluster has published its histogram

        // ---- block-level: rows of each bin in the whole block, and in the tiles in front of mine
        for (uint32_t b = t; b < P1; b += PB_THREADS) {
            uint32_t tot = 0, front = 0;
#pragma unroll
            for (uint32_t r = 0; r < (uint32_t)PB_CLUSTER; r++) {
                const uint32_t c = r == crank ? hist[b] : dsmem_ld_u32(hist + b, r);
                tot += c;
                if (r < crank) front += c;
            }
            s_tot[b] = tot;
            s_delta[b] = front;                                 // completed below
        }
        __syncthreads();
        if (warp == 0) {
            const uint32_t per = (P1 + 31) / 32;
            const uint32_t lo = min((uint32_t)lane * per, P1), hi = min(lo + per, P1);
            uint32_t sum = 0;
            for (uint32_t i = lo; i < hi; i++) sum += s_tot[i];
            uint32_t incl = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += y; }
            uint32_t run = incl - sum;
            const uint32_t base = a.out_base + pb_block_base(blk, mis);
            uint32_t *seg = a.seg + (size_t)blk * (P + 2);
            if (crank == 0 && lane == 0) a.block_base[blk] = base;
            for (uint32_t i = lo; i < hi; i++) {
                const uint32_t c = s_tot[i];
                s_delta[i] = base + run + s_delta[i] - s_off[i];
                if (crank == 0) {
                    seg[i] = run;
                    if (a.part_rows != nullptr && c) atomicAdd(a.part_rows + i, (unsigned long long)c);
                }
                run += c;
            }
            if (crank == 0 && lane == 31) seg[P1] = incl;      // rows of the block
        }
        __syncthreads();

        // ---- permutation: element (bin, rank) sits at s_off[bin] + offset of its warp + rank and remembers its source index
        if (have) {
#pragma unroll
            for (int q = 0; q < PB_ITEMS; q++) {
                const uint32_t bin = packed[q] >> 16, rank = packed[q] & 0xffffu;
                if (!inner && bin == P1) continue;
                const uint32_t src = ((uint32_t)(q / PER_VEC) * PB_THREADS + t) * PER_VEC + (q % PER_VEC);
#if B200J_PB_RANK == 2
                s_perm[s_off[bin] + rank] = (bin << 16) | src;
#else
                s_perm[s_off[bin] + my_wcnt[bin] + rank] = (bin << 16) | src;
#endif
            }
        }
        __syncthreads();

        // ---- write-out: consecutive threads -> consecutive elements of a bin's run -> consecutive addresses
        if (have) {
            const uint32_t total = inner ? (uint32_t)PB_TILE : s_scan[0];
            const uint32_t rid0 = vbase - mis + a.rid_base + a.out_base;
#pragma unroll 16
            for (uint32_t i = t; i < total; i += PB_THREADS) {
                const uint32_t pm = s_perm[i];
                const uint32_t p = pm >> 16, src = pm & 0xffffu;
                const uint32_t dst = s_delta[p] + i;
                out_keys[dst] = (OutT)s_stage[src];
                a.out_rid[dst] = rid0 + src;
            }
        }
        __syncthreads();
        fetch((blk + blk_step) * PB_CLUSTER + crank);
    }
    if (NARROW && __any_sync(0xffffffffu, wide_seen) && lane == 0) atomicExch(a.narrow_ovf, 1u);     // keys of rows outside the batch are zeros
    // a CTA's shared memory must outlive the peers' reads of it
    cluster_arrive();
    cluster_wait();
}

}  // namespace b200j
