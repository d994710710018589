// probe_dense.cuh — the dense probe with its key stream and its result stream moved off the load/store unit.
//
// Why: ncu of probe_dense_kernel (profiles/r1_d_ncu_full_bench_C2.txt) shows 1.46 G load sectors + 0.43 G store sectors
// through L1TEX per 1 B probe rows, and the kernel's 6.97 ms is exactly that count at the measured rate of ONE sector
// per clock per SM (scripts/ubench/gather.cu: 286 G random sectors/s = 0.98 / clk / SM). The table lookups themselves
// need 1.1 sectors per row; everything else (0.25 G sectors of key loads, 0.43 G of result stores, of which 0.18 G are
// stores split around pending lookups) is traffic that does not have to go through the LSU at all:
//   keys     every warp streams its own 1 KB key chunks (128 rows) into shared memory with bulk-async copies
//            (cp.async.bulk + mbarrier, UBLKCP in SASS), two chunks ahead of the one it is working on
//   results  build indices are staged in shared memory and leave with ONE bulk-async store per chunk
//            (cp.async.bulk.global.shared::cta, UBLKCP... .S2G), pending lookups included: a chunk is stored only after
//            the warp's queue has resolved every lookup of that chunk (ring of four staged chunks, FIFO queue)
// so the LSU is left with the random table sectors only. No block barrier anywhere: warps claim their own tiles.
#pragma once
#include "radix.cuh"
#include "probe.cuh"

namespace b200j {

constexpr int PW_WARPS  = 4;                      // warps per CTA (independent of each other)
constexpr int PW_CHUNK  = 128;                    // rows per chunk = PD_ITEMS rows per lane
#ifndef B200J_PW_TILE
#define B200J_PW_TILE 4096
#endif
constexpr int PW_TILE   = B200J_PW_TILE;          // rows per claim (one global atomic per warp)
constexpr int PW_CPT    = PW_TILE / PW_CHUNK;     // chunks per tile
constexpr int PW_KBUF   = 3;                      // key chunks in flight (lookahead 2)
constexpr int PW_RBUF   = 4;                      // staged result chunks
constexpr int PW_QCAP   = 160;                    // queue entries: < 32 carried over + at most 128 new
static_assert(PD_ITEMS * 32 == PW_CHUNK, "a chunk is one warp tile of the dense probe");

struct __align__(128) PWWarp {
    unsigned long long key[PW_KBUF][PW_CHUNK];    // 8 B per row reserved; 32-bit keys use the first half of a buffer
    unsigned long long res[PW_RBUF][PW_CHUNK];
    unsigned long long qkey[PW_QCAP];
    uint32_t           qloc[PW_QCAP];             // partition << 9 | result slot << 7 | row in chunk
    uint32_t           base[PW_RBUF];             // first row of the chunk staged in every result slot (lane 0 only)
    uint64_t           bar[PW_KBUF];
};

__device__ __forceinline__ void bulk_store(void *gdst, const void *ssrc, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(ssrc)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// One complete lookup (home bucket onwards) by a single lane: tail rows only.
__device__ __forceinline__ unsigned long long pw_lookup_one(const ProbeArgs &a, uint64_t key) {
    const uint64_t h = hash_one(a.hk, key);
    uint32_t base, nb;
    locate<true>(a.tv, (uint32_t)(h >> 32), base, nb);
    uint32_t b = home_bucket(key, nb);
    const Slot *p = key ? bucket_ptr(a.tv, base, b) : a.tv.zero_slot;
    for (;;) {
        unsigned long long k0, v0, k1, v1;
        load_bucket(p, k0, v0, k1, v1);
        if (k0 == key) return (uint32_t)(v0 >> 32) ? (unsigned long long)(uint32_t)v0 : ~0ULL;
        if (k1 == key) return (uint32_t)(v1 >> 32) ? (unsigned long long)(uint32_t)v1 : ~0ULL;
        if (k1 == 0ULL) return ~0ULL;
        b = next_bucket(b, nb);
        p = bucket_ptr(a.tv, base, b);
    }
}

template <int KW, bool MARK>
__global__ void __launch_bounds__(PW_WARPS * 32) probe_dense_tma_kernel(ProbeArgs a) {
    constexpr uint32_t FULL = 0xffffffffu, NONE = 0xffffffffu;
    constexpr int PER_VEC = 16 / KW;
    constexpr int NVEC = PD_ITEMS / PER_VEC;
    constexpr uint32_t KBYTES = (uint32_t)PW_CHUNK * KW;
    __shared__ PWWarp sh[PW_WARPS];
    PWWarp &w = sh[threadIdx.x >> 5];
    const int lane = threadIdx.x & 31;
    const uint32_t lt = (1u << lane) - 1u;
    const uint32_t n = (uint32_t)a.n;
    const uint32_t n_full = n & ~(uint32_t)(PW_CHUNK - 1);            // rows covered by whole chunks
    const uint32_t n_tiles = (uint32_t)(((uint64_t)n_full + PW_TILE - 1) / PW_TILE);
    const char *gkeys = reinterpret_cast<const char *>(a.keys);
    unsigned long long *gout = reinterpret_cast<unsigned long long *>(a.out_build);
    uint32_t misses = 0;

    // ---- the last n % 128 rows: first warp of the grid, plain loads and stores
    if (blockIdx.x == 0 && threadIdx.x < 32) {
        for (uint32_t r = n_full + lane; r < n; r += 32) {
            const uint64_t key = KW == 8 ? (uint64_t)reinterpret_cast<const unsigned long long *>(a.keys)[r]
                                         : (uint64_t)reinterpret_cast<const uint32_t *>(a.keys)[r];
            const unsigned long long res = pw_lookup_one(a, key);
            gout[r] = res;
            if (res == ~0ULL) misses++;
            else if (MARK) mark_visited(a.visited, (uint32_t)res);
        }
    }

    if (lane == 0) {
#pragma unroll
        for (int b = 0; b < PW_KBUF; b++) mbar_init(&w.bar[b], 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncwarp();

    // ---- this warp's sequence of chunks: tiles are claimed one ahead, chunk k lives in tile (k / PW_CPT) of the sequence
    uint32_t tile_a = 0, tile_b = NONE, claim_pending = 0;
    if (lane == 0) tile_a = atomicAdd(&a.ctr->next_tile, 1u);
    tile_a = __shfl_sync(FULL, tile_a, 0);
    uint32_t k = 0;
    auto chunk_row = [&](uint32_t kk) -> uint32_t {
        const uint32_t t = (kk / PW_CPT == k / PW_CPT) ? tile_a : tile_b;
        if (t >= n_tiles) return NONE;
        const uint32_t row = t * (uint32_t)PW_TILE + (kk % PW_CPT) * (uint32_t)PW_CHUNK;
        return row < n_full ? row : NONE;
    };
    auto issue_load = [&](uint32_t kk, uint32_t row) {             // lane 0; the buffer's last readers have passed a __syncwarp
        const uint32_t b = kk % PW_KBUF;
        fence_proxy_async();
        mbar_expect_tx(&w.bar[b], KBYTES);
        bulk_load(w.key[b], gkeys + (size_t)row * KW, KBYTES, &w.bar[b]);
    };

    // The stream is in partition order and a warp's chunks come in increasing row order, so the warp walks the partition
    // offsets once in its life. A chunk that lies inside one partition needs NO row hash: the partition's table geometry is
    // warp-uniform and the bucket comes from slot_hash(key). Only chunks straddling a boundary hash their rows.
    uint32_t p_cur = 0, p_base = 0, p_nb = 0;
    unsigned long long p_end = 0;         // first row past partition p_cur (0: not looked up yet)
    const uint32_t n_parts = 1u << a.tv.radix_bits;
    uint32_t q_head = 0, q_cnt = 0;       // FIFO ring of pending lookups
    uint32_t pend = 0;                    // pending lookups per result slot, 8 bits each
    uint32_t flushed = 0;                 // chunks [0, flushed) have left with a bulk store

    // oldest m (<= 32) queued lookups, one per lane: walk the probe sequence from the bucket after the home bucket
    auto drain_oldest = [&](uint32_t m) {
        if ((uint32_t)lane < m) {
            uint32_t e = q_head + lane;
            if (e >= PW_QCAP) e -= PW_QCAP;
            const uint64_t key = w.qkey[e];
            const uint32_t loc = w.qloc[e];
            const uint2 pe = __ldg(reinterpret_cast<const uint2 *>(a.tv.parts + (loc >> 9)));     // the entry remembers its partition
            const uint32_t base = pe.x, nb = pe.y;
            uint32_t b = next_bucket(home_bucket(key, nb), nb);
            unsigned long long res = ~0ULL;
            for (;;) {
                const uint32_t b2 = next_bucket(b, nb);
                unsigned long long k0, v0, k1, v1, k2, v2, k3, v3;
                load_bucket(bucket_ptr(a.tv, base, b), k0, v0, k1, v1);
                load_bucket(bucket_ptr(a.tv, base, b2), k2, v2, k3, v3);
                if (k0 == key) { res = (uint32_t)v0; break; }
                if (k1 == key) { res = (uint32_t)v1; break; }
                if (k1 == 0ULL) break;
                if (k2 == key) { res = (uint32_t)v2; break; }
                if (k3 == key) { res = (uint32_t)v3; break; }
                if (k3 == 0ULL) break;
                b = next_bucket(b2, nb);
            }
            w.res[(loc >> 7) & 3u][loc & 127u] = res;
            if (res == ~0ULL) misses++;
            else if (MARK) mark_visited(a.visited, (uint32_t)res);
        }
        __syncwarp();
        q_head += m;
        if (q_head >= PW_QCAP) q_head -= PW_QCAP;
        q_cnt -= m;
        // the queue is in chunk order: the drained entries belong to the oldest chunks that still have pending lookups
        uint32_t rem = m, f = flushed;
        while (rem) {
            const uint32_t sft = 8u * (f & 3u);
            const uint32_t d = min((pend >> sft) & 0xffu, rem);
            pend -= d << sft;
            rem -= d;
            f++;
        }
    };
    // chunks below `limit` are staged; store, in order, those without pending lookups
    auto flush_ready = [&](uint32_t limit) {
        while (flushed < limit && ((pend >> (8u * (flushed & 3u))) & 0xffu) == 0u) {
            fence_proxy_async();                      // every lane's staged results -> visible to the async proxy
            __syncwarp();
            if (lane == 0) {
                const uint32_t s = flushed & 3u;
                bulk_store(gout + w.base[s], w.res[s], (uint32_t)PW_CHUNK * 8u);
                bulk_commit();
            }
            flushed++;
        }
    };

    {
        const uint32_t r0 = chunk_row(0), r1 = chunk_row(1);
        if (lane == 0) {
            if (r0 != NONE) issue_load(0, r0);
            if (r1 != NONE) issue_load(1, r1);
        }
    }
    for (;; k++) {
        if (k != 0u && k % PW_CPT == 0u) { tile_a = tile_b; tile_b = NONE; }      // the warp moves on to the tile it claimed ahead
        const uint32_t rb = chunk_row(k);
        if (rb == NONE) break;
        if (k % PW_CPT == 0 && lane == 0) claim_pending = atomicAdd(&a.ctr->next_tile, 1u);     // answer needed 29 chunks from now
        if (k % PW_CPT == PW_CPT - 3) tile_b = __shfl_sync(FULL, claim_pending, 0);
        __syncwarp();                                  // chunk k-1's key reads are done: its buffer takes chunk k+2
        {
            const uint32_t r2 = chunk_row(k + 2);
            if (r2 != NONE && lane == 0) issue_load(k + 2, r2);
        }
        const uint32_t s = k & 3u;
        if (k >= PW_RBUF) {
            // result slot s still holds chunk k-4: it must have left, and its bulk store must have read the slot
            while (flushed + 3u < k) {
                drain_oldest(min(32u, q_cnt));
                flush_ready(k);
            }
            if (lane == 0) {
                const uint32_t newer = flushed - (k - 3u);         // bulk groups committed after chunk k-4's
                if (newer == 0u) bulk_wait_read<0>();
                else if (newer == 1u) bulk_wait_read<1>();
                else if (newer == 2u) bulk_wait_read<2>();
                else bulk_wait_read<3>();
            }
            __syncwarp();
        }
        if (lane == 0) w.base[s] = rb;
        mbar_wait(&w.bar[k % PW_KBUF], (k / PW_KBUF) & 1u);
        // ---- keys from shared memory: lane takes NVEC 16 B vectors, 32 vectors apart
        uint64_t key[PD_ITEMS];
        {
            const char *kb = reinterpret_cast<const char *>(w.key[k % PW_KBUF]);
#pragma unroll
            for (int v = 0; v < NVEC; v++) {
                if (KW == 8) {
                    const ulonglong2 x = *reinterpret_cast<const ulonglong2 *>(kb + ((uint32_t)v * 32u + lane) * 16u);
                    key[2 * v] = x.x; key[2 * v + 1] = x.y;
                } else {
                    const uint4 x = *reinterpret_cast<const uint4 *>(kb + ((uint32_t)v * 32u + lane) * 16u);
                    key[4 * v] = x.x; key[4 * v + 1] = x.y; key[4 * v + 2] = x.z; key[4 * v + 3] = x.w;
                }
            }
        }
        // ---- which partition(s) does this chunk lie in
        bool uniform = false;
        if (a.part_off != nullptr) {
            bool moved = false;
            if (p_end == 0) { p_end = __ldg(a.part_off + 1); moved = true; }           // first chunk of this warp: start at partition 0
            while (p_end <= (unsigned long long)rb && p_cur + 1u < n_parts) {
                p_cur++;
                p_end = __ldg(a.part_off + p_cur + 1);
                moved = true;
            }
            if (moved) {
                const uint2 pe = __ldg(reinterpret_cast<const uint2 *>(a.tv.parts + p_cur));
                p_base = pe.x;
                p_nb = pe.y;
            }
            uniform = (unsigned long long)rb + PW_CHUNK <= p_end;
        }
        // ---- one bucket load per row, all in flight
        unsigned long long k0[PD_ITEMS], v0[PD_ITEMS], k1[PD_ITEMS], v1[PD_ITEMS];
        uint32_t part[PD_ITEMS];
        if (uniform) {
#pragma unroll
            for (int q = 0; q < PD_ITEMS; q++) {
                part[q] = p_cur;
                const Slot *p = key[q] ? bucket_ptr(a.tv, p_base, home_bucket(key[q], p_nb)) : a.tv.zero_slot;
                load_bucket(p, k0[q], v0[q], k1[q], v1[q]);
            }
        } else {
#pragma unroll
            for (int q = 0; q < PD_ITEMS; q++) {
                const uint64_t h = hash_one(a.hk, key[q]);
                uint32_t base, nb;
                part[q] = (uint32_t)(h >> (64 - a.tv.radix_bits));
                locate<true>(a.tv, (uint32_t)(h >> 32), base, nb);
                const Slot *p = key[q] ? bucket_ptr(a.tv, base, home_bucket(key[q], nb)) : a.tv.zero_slot;
                load_bucket(p, k0[q], v0[q], k1[q], v1[q]);
            }
        }
        // ---- resolve; results (and placeholders of pending lookups) -> the chunk's result slot
        unsigned long long res[PD_ITEMS];
        bool pnd[PD_ITEMS];
        uint32_t n_pend = 0;
#pragma unroll
        for (int q = 0; q < PD_ITEMS; q++) {
            const bool hit0 = k0[q] == key[q], hit1 = k1[q] == key[q];
            const unsigned long long v = hit0 ? v0[q] : v1[q];
            const bool hit = (hit0 | hit1) && (uint32_t)(v >> 32) != 0u;      // count 0: the zero-key side bucket when key 0 is absent
            pnd[q] = !(hit0 | hit1) && k1[q] != 0ULL;
            res[q] = hit ? (unsigned long long)(uint32_t)v : ~0ULL;
            n_pend += pnd[q] ? 1u : 0u;
            if (!pnd[q]) {
                if (!hit) misses++;
                else if (MARK) mark_visited(a.visited, (uint32_t)v);
            }
        }
#pragma unroll
        for (int v = 0; v < NVEC; v++) {
            unsigned long long *o = &w.res[s][((uint32_t)v * 32u + lane) * PER_VEC];
#pragma unroll
            for (int j = 0; j < PER_VEC; j += 2)
                *reinterpret_cast<ulonglong2 *>(o + j) = make_ulonglong2(res[v * PER_VEC + j], res[v * PER_VEC + j + 1]);
        }
        // ---- queue the pending ones (warp-private FIFO: ballots, no atomics)
        const uint32_t b0 = __ballot_sync(FULL, n_pend & 1u), b1 = __ballot_sync(FULL, n_pend & 2u), b2 = __ballot_sync(FULL, n_pend & 4u);
        const uint32_t wtot = __popc(b0) + 2u * __popc(b1) + 4u * __popc(b2);
        if (wtot) {
            uint32_t pos = q_head + q_cnt + __popc(b0 & lt) + 2u * __popc(b1 & lt) + 4u * __popc(b2 & lt);
            if (pos >= PW_QCAP) pos -= PW_QCAP;
#pragma unroll
            for (int q = 0; q < PD_ITEMS; q++) {
                if (!pnd[q]) continue;
                w.qkey[pos] = key[q];
                w.qloc[pos] = (part[q] << 9) | (s << 7) | (((uint32_t)(q / PER_VEC) * 32u + lane) * PER_VEC + (q % PER_VEC));
                pos = pos + 1u == PW_QCAP ? 0u : pos + 1u;
            }
            q_cnt += wtot;
            pend += wtot << (8u * s);
        }
        __syncwarp();
        while (q_cnt >= 32u) drain_oldest(32u);
        flush_ready(k + 1u);
    }
    while (q_cnt) drain_oldest(min(32u, q_cnt));
    flush_ready(k);
    if (lane == 0) bulk_wait_all();
    const unsigned long long z = warp_sum_u32(misses);
    if (lane == 0 && z) atomicAdd(&a.ctr->build_null_out, z);
}

}  // namespace b200j
