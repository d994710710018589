// probe.cuh — the lean probe kernel (the dominant kernel of the join).
//
// Replaces, per probe batch, get_matching_indices[_with_probe] (src/shared/shared.rs:29-81) + Version10Lookup::get_iter
// (version10/lookup_implementation_3.rs:31-59) + equal_rows_arr (src/shared/datafusion_private.rs:40-80, fused: slots
// hold key VALUES) + the per-batch index algebra of the eight join types (datafusion_private.rs:85-185).
//
// Shape (why it looks like this — ncu of the first version showed 172 thread-instructions per row and 3.2
// serialised dependent L2 round trips per warp iteration for the ~10 % of lookups whose home bucket is full):
//   fast path   every thread hashes PF_ITEMS keys and issues PF_ITEMS independent 32 B bucket loads (one sector =
//               one bucket of two 16 B slots), then resolves each lookup from that single bucket: hit with one
//               build row, or provable miss (the bucket has a free slot). No loops, no dynamic register indexing.
//   slow queue  lookups that need a second bucket, and duplicate-key runs, are pushed to a shared-memory queue and
//               drained densely (one queued lookup per lane) once per super-tile, so stragglers never hold a warp
//               of finished lookups hostage.
//   staging     single-row results are compacted into shared memory (ballot positions, one shared atomic per warp)
//               and flushed with ONE global atomic per super-tile of PF_STAGE rows and fully coalesced stores.
//   scheduling  super-tiles are handed out by a global counter so the running CTAs sweep the partitioned probe
//               stream front to back: the one or two partitions' tables being probed stay in L2.
// Requires 16 B-aligned keys without a validity bitmap (always true for the radix layout, whose partition pass
// strips NULL keys; the direct layout falls back to probe_kernel in kernels.cuh otherwise).
#pragma once
#include "kernels.cuh"

namespace b200j {

constexpr int PF_THREADS = 256;
constexpr int PF_ITEMS   = 4;
constexpr int PF_TILE    = PF_THREADS * PF_ITEMS;     // 1024 rows
constexpr int PF_SUB     = 2;
constexpr int PF_STAGE   = PF_TILE * PF_SUB;          // 2048 rows per global atomic

struct PFShared {
    unsigned long long qkey[PF_STAGE];                // slow queue
    uint32_t           qrow[PF_STAGE];
    uint32_t           brow[PF_STAGE];                // staged single-row results (0xffffffff = NULL build side)
    uint32_t           prow[PF_STAGE];
    unsigned long long gbase;
    uint32_t           count, qcount, next, mcount;
};
constexpr uint32_t PF_SHORT = 8;                      // duplicate runs up to this length are staged like single rows

// Copies the duplicate-key runs of one warp's lanes (one run per lane, `mine` lanes only) to the output and/or marks
// them visited. Short runs are copied by their lane, long runs by the whole warp (coalesced).
template <int MODE, bool MARK>
__device__ __forceinline__ void emit_runs(const ProbeArgs &a, bool mine, uint32_t scnt, uint32_t srow, uint32_t prow, int lane) {
    constexpr uint32_t FULL = 0xffffffffu;
    constexpr bool emit = (MODE == PM_PAIRS || MODE == PM_OUTER);
    uint32_t off = 0;
    unsigned long long gbase = 0;
    if (emit) {
        const uint32_t mq = mine ? scnt : 0u;
        uint32_t incl = mq;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += y; }
        off = incl - mq;
        const uint32_t wtot = __shfl_sync(FULL, incl, 31);
        if (lane == 0) gbase = atomicAdd(&a.ctr->out_count, (unsigned long long)wtot);
        gbase = __shfl_sync(FULL, gbase, 0);
    }
    if (mine && scnt < 32u) {
        const unsigned long long pos = gbase + off;
        for (uint32_t k = 0; k < scnt; k++) {
            const uint32_t br = __ldcg(a.tv.rowids + srow + k);
            if (emit && pos + k < a.out_cap) {
                a.out_build[pos + k] = (uint64_t)br;
                a.out_probe[pos + k] = prow;
            }
            if (MARK) mark_visited(a.visited, br);
        }
    }
    uint32_t pend = __ballot_sync(FULL, mine && scnt >= 32u);
    while (pend) {
        const int src = __ffs(pend) - 1;
        pend &= pend - 1;
        const unsigned long long pos = __shfl_sync(FULL, gbase + off, src);
        const uint32_t cnt = __shfl_sync(FULL, scnt, src);
        const uint32_t rb = __shfl_sync(FULL, srow, src);
        const uint32_t pr = __shfl_sync(FULL, prow, src);
        copy_long_run<emit, MARK>(a, pos, cnt, rb, pr, lane);
    }
}

template <int KW, int MODE, bool MARK, bool RADIX>
__global__ void __launch_bounds__(PF_THREADS) probe_fast_kernel(ProbeArgs a) {
    constexpr uint32_t FULL = 0xffffffffu;
    constexpr int PER_VEC = 16 / KW;
    constexpr int NVEC = PF_ITEMS / PER_VEC;
    __shared__ PFShared sh;
    const int t = threadIdx.x, lane = t & 31;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const uint32_t n = (uint32_t)a.n;
    const uint32_t n_tiles = (n + PF_TILE - 1) / PF_TILE;
    const uint32_t n_super = (n_tiles + PF_SUB - 1) / PF_SUB;
    uint32_t null_build_out = 0;

    if (t == 0) { sh.count = 0; sh.qcount = 0; sh.mcount = 0; sh.next = atomicAdd(&a.ctr->next_tile, 1u); }
    __syncthreads();
    for (;;) {
        const uint32_t sup = sh.next;
        if (sup >= n_super) break;
#pragma unroll 1
        for (int sub = 0; sub < PF_SUB; sub++) {
            const uint32_t tb = (sup * PF_SUB + sub) * (uint32_t)PF_TILE;
            if (tb >= n) break;
            // ---- keys + probe row ids of this thread's items
            uint64_t key[PF_ITEMS];
            uint32_t prow[PF_ITEMS];
            bool inb[PF_ITEMS];
#pragma unroll
            for (int v = 0; v < NVEC; v++) {
                const uint32_t r0 = tb + ((uint32_t)v * PF_THREADS + t) * PER_VEC;
                if (KW == 8) {
                    const unsigned long long *k64 = reinterpret_cast<const unsigned long long *>(a.keys);
                    inb[2 * v] = r0 < n;
                    inb[2 * v + 1] = r0 + 1 < n;
                    if (inb[2 * v + 1]) {
                        const ulonglong2 x = __ldcs(reinterpret_cast<const ulonglong2 *>(k64 + r0));
                        key[2 * v] = x.x; key[2 * v + 1] = x.y;
                        if (RADIX) { const uint2 r = __ldcs(reinterpret_cast<const uint2 *>(a.rowid + r0)); prow[2 * v] = r.x; prow[2 * v + 1] = r.y; }
                    } else {
                        key[2 * v] = inb[2 * v] ? __ldcs(k64 + r0) : 0ULL;
                        key[2 * v + 1] = 0ULL;
                        if (RADIX) { prow[2 * v] = inb[2 * v] ? __ldcs(a.rowid + r0) : 0u; prow[2 * v + 1] = 0u; }
                    }
                    if (!RADIX) { prow[2 * v] = r0; prow[2 * v + 1] = r0 + 1; }
                } else {
                    const uint32_t *k32 = reinterpret_cast<const uint32_t *>(a.keys);
#pragma unroll
                    for (int q = 0; q < 4; q++) inb[4 * v + q] = r0 + q < n;
                    if (inb[4 * v + 3]) {
                        const uint4 x = __ldcs(reinterpret_cast<const uint4 *>(k32 + r0));
                        key[4 * v] = x.x; key[4 * v + 1] = x.y; key[4 * v + 2] = x.z; key[4 * v + 3] = x.w;
                        if (RADIX) {
                            const uint4 r = __ldcs(reinterpret_cast<const uint4 *>(a.rowid + r0));
                            prow[4 * v] = r.x; prow[4 * v + 1] = r.y; prow[4 * v + 2] = r.z; prow[4 * v + 3] = r.w;
                        }
                    } else {
#pragma unroll
                        for (int q = 0; q < 4; q++) {
                            key[4 * v + q] = inb[4 * v + q] ? (uint64_t)__ldcs(k32 + r0 + q) : 0ULL;
                            if (RADIX) prow[4 * v + q] = inb[4 * v + q] ? __ldcs(a.rowid + r0 + q) : 0u;
                        }
                    }
                    if (!RADIX) {
#pragma unroll
                        for (int q = 0; q < 4; q++) prow[4 * v + q] = r0 + q;
                    }
                }
            }
            // ---- hash + one bucket load per item, all in flight before any is consumed
            unsigned long long k0[PF_ITEMS], v0[PF_ITEMS], k1[PF_ITEMS], v1[PF_ITEMS];
#pragma unroll
            for (int q = 0; q < PF_ITEMS; q++) {
                const uint64_t h = hash_one(a.hk, key[q]);
                uint32_t base, nb;
                locate<RADIX>(a.tv, (uint32_t)(h >> 32), base, nb);
                const Slot *p = key[q] ? bucket_ptr(a.tv, base, home_bucket(h, nb)) : a.tv.zero_slot;
                load_bucket(p, k0[q], v0[q], k1[q], v1[q], (!RADIX && a.tv.l1 != 0u));
            }
            // ---- resolve from that bucket: hit, provable miss (bucket has a free slot; slots fill in order, so a free
            //      slot shows in the odd one), or "needs the slow path" (bucket full without the key, or a run)
            bool one[PF_ITEMS], slow[PF_ITEMS];
            uint32_t srow[PF_ITEMS], n_one = 0, n_slow = 0;
#pragma unroll
            for (int q = 0; q < PF_ITEMS; q++) {
                const bool hit0 = k0[q] == key[q], hit1 = k1[q] == key[q];
                const unsigned long long v = hit0 ? v0[q] : (hit1 ? v1[q] : 0ULL);
                const uint32_t cnt = (uint32_t)(v >> 32) & COUNT_MASK;
                const bool resolved = hit0 | hit1 | (k1[q] == 0ULL) | !bucket_overflowed(v0[q]);      // home bucket: see SLOT_OVF
                srow[q] = (uint32_t)v;
                bool o, s;
                if (MODE == PM_PAIRS) { o = cnt == 1u; s = !resolved || cnt > 1u; }
                else if (MODE == PM_OUTER) { o = resolved && cnt <= 1u; s = !resolved || cnt > 1u; if (cnt == 0u) srow[q] = 0xffffffffu; }
                else if (MODE == PM_SEMI) { o = cnt != 0u; s = !resolved; srow[q] = 0xffffffffu; }
                else if (MODE == PM_ANTI) { o = resolved && cnt == 0u; s = !resolved; srow[q] = 0xffffffffu; }
                else { o = false; s = !resolved || cnt > 1u; if (cnt == 1u && inb[q]) mark_visited(a.visited, (uint32_t)v); }
                one[q] = o && inb[q];
                slow[q] = s && inb[q];
                n_one += one[q] ? 1u : 0u;
                n_slow += slow[q] ? 1u : 0u;
                if (MARK && MODE != PM_MARK && one[q] && cnt == 1u) mark_visited(a.visited, (uint32_t)v);
            }
            // ---- single rows -> shared-memory stage. Every lane writes its rows back to back; the lane's base is the
            //      warp-exclusive prefix of n_one (0..4), taken from three ballots over its bits.
            if (MODE != PM_MARK) {
                const uint32_t b0 = __ballot_sync(FULL, n_one & 1u), b1 = __ballot_sync(FULL, n_one & 2u), b2 = __ballot_sync(FULL, n_one & 4u);
                const uint32_t wtot = __popc(b0) + 2u * __popc(b1) + 4u * __popc(b2);
                uint32_t pos = __popc(b0 & lt_mask) + 2u * __popc(b1 & lt_mask) + 4u * __popc(b2 & lt_mask);
                uint32_t wbase = 0;
                if (lane == 0 && wtot) wbase = atomicAdd(&sh.count, wtot);
                pos += __shfl_sync(FULL, wbase, 0);
#pragma unroll
                for (int q = 0; q < PF_ITEMS; q++) {
                    if (!one[q]) continue;
                    sh.brow[pos] = srow[q];
                    sh.prow[pos] = prow[q];
                    pos++;
                    if (MODE != PM_PAIRS && srow[q] == 0xffffffffu) null_build_out++;
                }
            }
            // ---- everything else -> slow queue (one reservation per lane; ptxas aggregates it per warp)
            if (n_slow) {
                uint32_t pos = atomicAdd(&sh.qcount, n_slow);
#pragma unroll
                for (int q = 0; q < PF_ITEMS; q++) {
                    if (!slow[q]) continue;
                    sh.qkey[pos] = key[q];
                    sh.qrow[pos] = prow[q];
                    pos++;
                }
            }
        }
        __syncthreads();
        // ---- drain the slow queue: one queued lookup per lane, full probe-sequence walk
        const uint32_t qn = sh.qcount;
        for (uint32_t i0 = 0; i0 < qn; i0 += PF_THREADS) {
            const uint32_t i = i0 + t;
            const bool act = i < qn;
            const uint64_t key = act ? sh.qkey[i] : 0ULL;
            const uint32_t prow = act ? sh.qrow[i] : 0u;
            __syncthreads();          // every entry of this round is in registers: the duplicate-run list below may reuse the queue
            uint32_t srow = 0, scnt = 0;
            if (act) {
                const uint64_t h = hash_one(a.hk, key);
                uint32_t base, nb;
                locate<RADIX>(a.tv, (uint32_t)(h >> 32), base, nb);
                uint32_t idx = home_bucket(h, nb);
                const Slot *p = key ? bucket_ptr(a.tv, base, idx) : a.tv.zero_slot;
                for (;;) {
                    unsigned long long k0, v0, k1, v1;
                    load_bucket(p, k0, v0, k1, v1, (!RADIX && a.tv.l1 != 0u));
                    const bool hit0 = k0 == key, hit1 = k1 == key;
                    if (hit0 | hit1 | (k1 == 0ULL)) {
                        const unsigned long long v = hit0 ? v0 : (hit1 ? v1 : 0ULL);
                        srow = (uint32_t)v;
                        scnt = (uint32_t)(v >> 32) & COUNT_MASK;
                        break;
                    }
                    idx = next_bucket(idx, nb);
                    p = bucket_ptr(a.tv, base, idx);
                }
            }
            const bool found = act && scnt != 0u;
            bool o = false, multi = false;
            uint32_t sr = srow;
            if (MODE == PM_PAIRS) { o = found && scnt == 1u; multi = found && scnt > 1u; }
            else if (MODE == PM_OUTER) { o = act && scnt <= 1u; multi = found && scnt > 1u; if (!found) sr = 0xffffffffu; }
            else if (MODE == PM_SEMI) { o = found; sr = 0xffffffffu; }
            else if (MODE == PM_ANTI) { o = act && !found; sr = 0xffffffffu; }
            else { multi = found && scnt > 1u; if (found && scnt == 1u) mark_visited(a.visited, srow); }
            if (MODE != PM_MARK) {
                const uint32_t b = __ballot_sync(FULL, o);
                uint32_t wbase = 0;
                if (lane == 0 && b) wbase = atomicAdd(&sh.count, (uint32_t)__popc(b));
                wbase = __shfl_sync(FULL, wbase, 0);
                if (o) {
                    const uint32_t pos = wbase + __popc(b & lt_mask);
                    sh.brow[pos] = sr;
                    sh.prow[pos] = prow;
                    if (MODE != PM_PAIRS && sr == 0xffffffffu) null_build_out++;
                    if (MARK && sr != 0xffffffffu) mark_visited(a.visited, sr);
                }
            }
            if (MODE == PM_PAIRS || MODE == PM_OUTER || MARK) {
                // duplicate runs wait until every single row of the super-tile is staged: (run start, length, probe row) goes
                // back into the queue arrays, compacted (position < number of entries already read, see the barrier above)
                const uint32_t bm = __ballot_sync(FULL, multi);
                uint32_t mbase = 0;
                if (lane == 0 && bm) mbase = atomicAdd(&sh.mcount, (uint32_t)__popc(bm));
                mbase = __shfl_sync(FULL, mbase, 0);
                if (multi) {
                    const uint32_t j = mbase + __popc(bm & lt_mask);
                    sh.qkey[j] = (unsigned long long)srow | ((unsigned long long)scnt << 32);
                    sh.qrow[j] = prow;
                }
            }
        }
        __syncthreads();
        // ---- duplicate runs. Short ones are staged like single rows while the stage has room (their pairs then leave in the
        //      coalesced flush: a per-lane copy straight to global memory costs two L2 sectors per output row); long ones, and
        //      whatever does not fit, reserve their own output range per warp (emit_runs).
        if (MODE == PM_PAIRS || MODE == PM_OUTER || MARK) {
            const uint32_t mn = sh.mcount;
            for (uint32_t j0 = 0; j0 < mn; j0 += PF_THREADS) {
                const uint32_t j = j0 + t;
                const bool act = j < mn;
                const unsigned long long v = act ? sh.qkey[j] : 0ULL;
                const uint32_t srow = (uint32_t)v, scnt = (uint32_t)(v >> 32);
                const uint32_t prow = act ? sh.qrow[j] : 0u;
                bool staged = false;
                if (MODE == PM_PAIRS || MODE == PM_OUTER) {
                    const bool shortrun = act && scnt <= PF_SHORT;
                    const uint32_t need = shortrun ? scnt : 0u;
                    uint32_t incl = need;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += y; }
                    const uint32_t wtot = __shfl_sync(FULL, incl, 31);
                    uint32_t wbase = 0xffffffffu;
                    if (lane == 0 && wtot) {
                        uint32_t old = *reinterpret_cast<volatile uint32_t *>(&sh.count);
                        while (old + wtot <= (uint32_t)PF_STAGE) {
                            const uint32_t prev = atomicCAS(&sh.count, old, old + wtot);
                            if (prev == old) { wbase = old; break; }
                            old = prev;
                        }
                    }
                    wbase = __shfl_sync(FULL, wbase, 0);
                    staged = shortrun && wbase != 0xffffffffu;
                    if (staged) {
                        const uint32_t pos = wbase + incl - need;
                        for (uint32_t k = 0; k < scnt; k++) {
                            const uint32_t br = __ldcg(a.tv.rowids + srow + k);
                            sh.brow[pos + k] = br;
                            sh.prow[pos + k] = prow;
                            if (MARK) mark_visited(a.visited, br);
                        }
                    }
                }
                const bool rest = act && !staged;
                if (__any_sync(FULL, rest)) emit_runs<MODE, MARK>(a, rest, scnt, srow, prow, lane);
            }
        }
        __syncthreads();
        // ---- flush: one global atomic per super-tile, then fully coalesced streaming stores
        const uint32_t cnt = sh.count;
        if (t == 0) sh.gbase = cnt ? atomicAdd(&a.ctr->out_count, (unsigned long long)cnt) : 0ULL;
        __syncthreads();
        const unsigned long long gbase = sh.gbase;
        if (gbase + cnt <= a.out_cap) {
            if (a.out_build)
                for (uint32_t i = t; i < cnt; i += PF_THREADS) {
                    const uint32_t br = sh.brow[i];
                    __stcs(reinterpret_cast<unsigned long long *>(a.out_build) + gbase + i, br == 0xffffffffu ? ~0ULL : (unsigned long long)br);
                }
            for (uint32_t i = t; i < cnt; i += PF_THREADS) __stcs(a.out_probe + gbase + i, sh.prow[i]);
        } else {
            for (uint32_t i = t; i < cnt; i += PF_THREADS) {
                if (gbase + i >= a.out_cap) break;
                const uint32_t br = sh.brow[i];
                if (a.out_build) a.out_build[gbase + i] = br == 0xffffffffu ? ~0ULL : (uint64_t)br;
                a.out_probe[gbase + i] = sh.prow[i];
            }
        }
        __syncthreads();
        if (t == 0) { sh.count = 0; sh.qcount = 0; sh.mcount = 0; sh.next = atomicAdd(&a.ctr->next_tile, 1u); }
        __syncthreads();
    }
    if (MODE == PM_OUTER || MODE == PM_SEMI || MODE == PM_ANTI) {
        const unsigned long long z = warp_sum_u32(null_build_out);
        if ((t & 31) == 0 && z) atomicAdd(&a.ctr->build_null_out, z);
    }
}


// =====================================================================================================================
// Dense probe (build keys unique): the output is NOT compacted. RADIX = false is the same kernel over the caller's own key
// buffer (direct layout, 16 B-aligned keys without a validity bitmap); it also writes the probe-index column (row i -> i).
// Radix layout: Row i of the partitioned probe stream
// writes out_build[i] = its build row or NULL; the partition pass's row-id array IS the probe-index column of the
// result, so this kernel neither reads nor writes it: 8 B key in, 8 B build index out per probe row (16 B/row of HBM
// traffic instead of 24 B), no shared-memory staging, no block barrier on the hot path, no output atomics.
// Inner / Left joins whose batch had misses get one compaction pass afterwards (compact_pairs_kernel); Right / Full
// want the NULL rows anyway. A foreign-key probe (every probe row matches once) is exactly dense.
//
// Lookups whose home bucket is full without the key (~10 % at load factor 0.5) go to a warp-private shared-memory
// queue and are drained 32 at a time — one per lane, two buckets fetched per round — so the warp's other 31 lanes
// never wait for one lane's probe sequence.
// =====================================================================================================================
constexpr int PD_THREADS = 256;
constexpr int PD_WARPS   = PD_THREADS / 32;
constexpr int PD_ITEMS   = 4;
constexpr int PD_WTILE   = 32 * PD_ITEMS;             // rows per warp iteration
#ifndef B200J_PD_CHUNK
#define B200J_PD_CHUNK 1024
#endif
constexpr int PD_CHUNK   = B200J_PD_CHUNK;             // rows per WARP claim (one global atomic); a multiple of PD_WTILE
#ifndef B200J_PD_MIN_CTAS
#define B200J_PD_MIN_CTAS 1
#endif
constexpr int PD_MIN_CTAS = B200J_PD_MIN_CTAS;        // measured with 5 (48 registers, no spills, 8-byte keys behind the radix pass): C2 probe kernel 7.04 -> 6.95 ms at load factor 0.5; 6 spills: 9.8 ms. Default stays 1 until the GPU suite has run with 5 (r1_l ended without a box)
constexpr int PD_QCAP    = 32 + PD_WTILE;             // a queue holds < 32 entries between iterations

struct PDShared {
    unsigned long long qkey[PD_WARPS][PD_QCAP];
    uint32_t           qidx[PD_WARPS][PD_QCAP];
};
static_assert(PD_CHUNK % PD_WTILE == 0, "a warp's claim is a whole number of its 128-row iterations");

#ifndef B200J_DRAIN_PAIR
#define B200J_DRAIN_PAIR 1      // a drained lookup fetches two consecutive buckets per round (0: one)
#endif
template <bool MARK, bool RADIX, bool L1, bool SLOT = false>
__device__ __forceinline__ void dense_drain(const ProbeArgs &a, unsigned long long *qkey, uint32_t *qidx, uint32_t first, uint32_t cnt,
                                            int lane, uint32_t &misses, uint32_t *out_slot = nullptr) {
    // entries [first, first + cnt) of the warp's queue, one per lane (cnt <= 32)
    const bool act = (uint32_t)lane < cnt;
    const uint64_t key = act ? qkey[first + lane] : 0ULL;
    const uint32_t oi = act ? qidx[first + lane] : 0u;
    if (act) {
        const uint64_t h = hash_one(a.hk, key);
        uint32_t base, nb;
        locate<RADIX>(a.tv, (uint32_t)(h >> 32), base, nb);
        uint32_t b = next_bucket(home_bucket(h, nb), nb);          // the home bucket was full without the key
        unsigned long long res = ~0ULL;
        uint32_t slot = 0xffffffffu;                               // SLOT: index of the slot that holds the key
#if B200J_DRAIN_PAIR
        for (;;) {
            const uint32_t b2 = next_bucket(b, nb);
            unsigned long long k0, v0, k1, v1, k2, v2, k3, v3;
            load_bucket(bucket_ptr(a.tv, base, b), k0, v0, k1, v1, L1);
            load_bucket(bucket_ptr(a.tv, base, b2), k2, v2, k3, v3, L1);
            if (k0 == key) { res = (uint32_t)v0; slot = 2u * (base + b); break; }
            if (k1 == key) { res = (uint32_t)v1; slot = 2u * (base + b) + 1u; break; }
            if (k1 == 0ULL) break;
            if (k2 == key) { res = (uint32_t)v2; slot = 2u * (base + b2); break; }
            if (k3 == key) { res = (uint32_t)v3; slot = 2u * (base + b2) + 1u; break; }
            if (k3 == 0ULL) break;
            b = next_bucket(b2, nb);
        }
#else
        for (;;) {
            unsigned long long k0, v0, k1, v1;
            load_bucket(bucket_ptr(a.tv, base, b), k0, v0, k1, v1, L1);
            if (k0 == key) { res = (uint32_t)v0; slot = 2u * (base + b); break; }
            if (k1 == key) { res = (uint32_t)v1; slot = 2u * (base + b) + 1u; break; }
            if (k1 == 0ULL) break;
            b = next_bucket(b, nb);
        }
#endif
        a.out_build[oi] = res;
        if (SLOT) out_slot[oi] = slot;
        if (res == ~0ULL) misses++;
        else if (MARK) mark_visited(a.visited, (uint32_t)res);
    }
}

// L1: bucket loads allocate in L1 -- a small direct table, or a probe batch with heavy-hitter keys (every SM otherwise asks
// the L2 for the same few sectors); off for uniform keys over a big table, where an L1 line is never reused.
template <int KW, bool MARK, bool RADIX, bool L1>
__global__ void __launch_bounds__(PD_THREADS, (KW == 8 && RADIX) ? PD_MIN_CTAS : 1) probe_dense_kernel(ProbeArgs a) {
    constexpr int PER_VEC = 16 / KW;
    constexpr int NVEC = PD_ITEMS / PER_VEC;
    constexpr uint32_t FULL = 0xffffffffu;
    __shared__ PDShared sh;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const uint32_t n = (uint32_t)a.n;
    unsigned long long *qkey = sh.qkey[warp];
    uint32_t *qidx = sh.qidx[warp];
    uint32_t qn = 0;                                      // warp-uniform
    uint32_t misses = 0;
    // Work is claimed per WARP in chunks of PD_CHUNK rows from one global counter (no block barrier anywhere in this kernel).
    // The rows in flight over the whole GPU are then a window of (resident warps x PD_CHUNK) consecutive rows of the
    // partitioned stream: measured with scripts/ubench/probe_bounds.cu, the probe runs at 203 G rows/s when every lookup hits
    // L2 and at 147 G rows/s when the window spans 2.3 partitions' tables (the CTA-wide 16 K-row claims of round 1: 9.7 M rows
    // in flight against 7.8 M rows per partition at C2). The next claim is issued before the current chunk is processed.
    uint32_t claim = 0;
    if (lane == 0) claim = atomicAdd(&a.ctr->next_tile, 1u);
    for (;;) {
        const uint32_t cur = __shfl_sync(FULL, claim, 0);
        if ((uint64_t)cur * PD_CHUNK >= (uint64_t)n) break;
        if (lane == 0) claim = atomicAdd(&a.ctr->next_tile, 1u);
        const uint32_t s0 = cur * (uint32_t)PD_CHUNK;
        const uint32_t s1 = (uint32_t)min((uint64_t)s0 + PD_CHUNK, (uint64_t)n);
        for (uint32_t wb = s0; wb < s1; wb += PD_WTILE) {
            // ---- keys: lane takes NVEC vectors, 32 vectors apart
            uint64_t key[PD_ITEMS];
            uint32_t idx0[NVEC];
            bool inb[PD_ITEMS];
#pragma unroll
            for (int v = 0; v < NVEC; v++) {
                const uint32_t r0 = wb + ((uint32_t)v * 32u + lane) * PER_VEC;
                idx0[v] = r0;
                if (KW == 8) {
                    const unsigned long long *k64 = reinterpret_cast<const unsigned long long *>(a.keys);
                    inb[2 * v] = r0 < s1;
                    inb[2 * v + 1] = r0 + 1 < s1;
                    if (inb[2 * v + 1]) {
                        const ulonglong2 x = __ldcs(reinterpret_cast<const ulonglong2 *>(k64 + r0));
                        key[2 * v] = x.x; key[2 * v + 1] = x.y;
                    } else {
                        key[2 * v] = inb[2 * v] ? __ldcs(k64 + r0) : 0ULL;
                        key[2 * v + 1] = 0ULL;
                    }
                } else {
                    const uint32_t *k32 = reinterpret_cast<const uint32_t *>(a.keys);
#pragma unroll
                    for (int q = 0; q < 4; q++) inb[4 * v + q] = r0 + q < s1;
                    if (inb[4 * v + 3]) {
                        const uint4 x = __ldcs(reinterpret_cast<const uint4 *>(k32 + r0));
                        key[4 * v] = x.x; key[4 * v + 1] = x.y; key[4 * v + 2] = x.z; key[4 * v + 3] = x.w;
                    } else {
#pragma unroll
                        for (int q = 0; q < 4; q++) key[4 * v + q] = inb[4 * v + q] ? (uint64_t)__ldcs(k32 + r0 + q) : 0ULL;
                    }
                }
            }
            // ---- one bucket load per item, all in flight
            unsigned long long k0[PD_ITEMS], v0[PD_ITEMS], k1[PD_ITEMS], v1[PD_ITEMS];
#pragma unroll
            for (int q = 0; q < PD_ITEMS; q++) {
                const uint64_t h = hash_one(a.hk, key[q]);
                uint32_t base, nb;
                locate<RADIX>(a.tv, (uint32_t)(h >> 32), base, nb);
                const Slot *p = key[q] ? bucket_ptr(a.tv, base, home_bucket(h, nb)) : a.tv.zero_slot;
                load_bucket(p, k0[q], v0[q], k1[q], v1[q], L1);
            }
            // ---- resolve; unresolved lookups go to the warp's queue
            unsigned long long res[PD_ITEMS];
            bool pend[PD_ITEMS];
            uint32_t n_pend = 0;
#pragma unroll
            for (int q = 0; q < PD_ITEMS; q++) {
                const bool hit0 = k0[q] == key[q], hit1 = k1[q] == key[q];
                const unsigned long long v = hit0 ? v0[q] : v1[q];
                const bool hit = (hit0 | hit1) && ((uint32_t)(v >> 32) & COUNT_MASK) != 0u;      // count 0: the zero-key side bucket when key 0 is absent
                pend[q] = inb[q] && !(hit0 | hit1) && k1[q] != 0ULL && bucket_overflowed(v0[q]);   // full home bucket that never overflowed: a miss
                res[q] = hit ? (unsigned long long)(uint32_t)v : ~0ULL;
                n_pend += pend[q] ? 1u : 0u;
                if (inb[q] && !pend[q]) {
                    if (!hit) misses++;
                    else if (MARK) mark_visited(a.visited, (uint32_t)v);
                }
            }
            // ---- dense store: 16 B per pair of rows when neither is pending
#pragma unroll
            for (int v = 0; v < NVEC; v++) {
                if (KW == 8) {
                    unsigned long long *o = reinterpret_cast<unsigned long long *>(a.out_build) + idx0[v];
                    if (inb[2 * v + 1] && !pend[2 * v] && !pend[2 * v + 1]) {
                        __stcs(reinterpret_cast<ulonglong2 *>(o), make_ulonglong2(res[2 * v], res[2 * v + 1]));
                    } else {
                        if (inb[2 * v] && !pend[2 * v]) __stcs(o, res[2 * v]);
                        if (inb[2 * v + 1] && !pend[2 * v + 1]) __stcs(o + 1, res[2 * v + 1]);
                    }
                } else {
                    unsigned long long *o = reinterpret_cast<unsigned long long *>(a.out_build) + idx0[v];
                    const bool all = inb[4 * v + 3] && !pend[4 * v] && !pend[4 * v + 1] && !pend[4 * v + 2] && !pend[4 * v + 3];
                    if (all) {
                        __stcs(reinterpret_cast<ulonglong2 *>(o), make_ulonglong2(res[4 * v], res[4 * v + 1]));
                        __stcs(reinterpret_cast<ulonglong2 *>(o) + 1, make_ulonglong2(res[4 * v + 2], res[4 * v + 3]));
                    } else {
#pragma unroll
                        for (int q = 0; q < 4; q++)
                            if (inb[4 * v + q] && !pend[4 * v + q]) __stcs(o + q, res[4 * v + q]);
                    }
                }
            }
            // ---- direct layout: no partition pass wrote the probe-index column, so it is written here (row i -> i)
            if (!RADIX) {
#pragma unroll
                for (int v = 0; v < NVEC; v++) {
                    uint32_t *o = a.out_probe + idx0[v];
                    if (KW == 8) {
                        if (inb[2 * v + 1]) __stcs(reinterpret_cast<uint2 *>(o), make_uint2(idx0[v], idx0[v] + 1u));
                        else if (inb[2 * v]) __stcs(o, idx0[v]);
                    } else {
                        if (inb[4 * v + 3]) __stcs(reinterpret_cast<uint4 *>(o), make_uint4(idx0[v], idx0[v] + 1u, idx0[v] + 2u, idx0[v] + 3u));
                        else {
#pragma unroll
                            for (int q = 0; q < 4; q++)
                                if (inb[4 * v + q]) __stcs(o + q, idx0[v] + (uint32_t)q);
                        }
                    }
                }
            }
            // ---- queue the pending ones (warp-private: ballots, no atomics)
            const uint32_t b0 = __ballot_sync(FULL, n_pend & 1u), b1 = __ballot_sync(FULL, n_pend & 2u), b2 = __ballot_sync(FULL, n_pend & 4u);
            const uint32_t wtot = __popc(b0) + 2u * __popc(b1) + 4u * __popc(b2);
            if (wtot) {
                const uint32_t lt = (1u << lane) - 1u;
                uint32_t pos = qn + __popc(b0 & lt) + 2u * __popc(b1 & lt) + 4u * __popc(b2 & lt);
#pragma unroll
                for (int q = 0; q < PD_ITEMS; q++) {
                    if (!pend[q]) continue;
                    qkey[pos] = key[q];
                    qidx[pos] = idx0[q / PER_VEC] + (q % PER_VEC);
                    pos++;
                }
                qn += wtot;
                __syncwarp();
                while (qn >= 32u) {
                    qn -= 32u;
                    dense_drain<MARK, RADIX, L1>(a, qkey, qidx, qn, 32u, lane, misses);
                    __syncwarp();
                }
            }
        }
    }
    if (qn) dense_drain<MARK, RADIX, L1>(a, qkey, qidx, 0u, qn, lane, misses);
    const unsigned long long z = warp_sum_u32(misses);
    if (lane == 0 && z) atomicAdd(&a.ctr->build_null_out, z);
}

// =====================================================================================================================
// Dense probe over the BLOCK-LOCALLY partitioned stream of radix_block.cuh (round 2). Unit u = (partition p, block b), p major:
// the rows of partition p inside block b are the segment [seg[b][p], seg[b][p + 1]) of the block's PB_BLOCK rows. Warps claim
// SEG_UNITS consecutive units from one global counter, so the whole GPU sweeps partition 0's segments, then partition 1's ...
// and the table of the partition being probed stays in L2 (the window of rows in flight is resident warps x ~1 K rows). Lane j
// of the warp fetches the bounds of unit j of the claim; the warp then walks the segments 128 rows at a time: 8 B key loads
// and 8 B result stores, 32 lanes contiguous, whatever the segment's alignment. The partition is known from the unit, not
// from the hash: no per-row descriptor load. p == P are the NULL-key rows: they never match (src/lib.rs:149-193) and get the
// NULL build index without a lookup. Pending lookups (home bucket full without the key) go through the same warp-private
// queue as in probe_dense_kernel.
// =====================================================================================================================
#ifndef B200J_SEG_FLAT
#define B200J_SEG_FLAT 0       // 1: the claim's segments walked as one virtual stream with key prefetch (measured slower: 7.6 vs 7.2 ms at C2)
#endif
#ifndef B200J_SEG_UNITS
#define B200J_SEG_UNITS 4
#endif
constexpr int SEG_UNITS = B200J_SEG_UNITS;             // units per warp claim (a power of two <= 32)
static_assert(SEG_UNITS >= 2 && SEG_UNITS <= 32 && (SEG_UNITS & (SEG_UNITS - 1)) == 0, "SEG_UNITS");
#ifndef B200J_SEG_MIN_CTAS
#define B200J_SEG_MIN_CTAS 4
#endif

struct SegArgs {
    const uint32_t *seg;          // [n_blocks][P + 2]
    const uint32_t *block_base;   // [n_blocks] first row of every block in the partitioned arrays
    uint32_t        n_blocks;
    uint32_t        n_units;      // n_blocks * P, or n_blocks * (P + 1) when the batch has NULL keys
    uint32_t       *out_slot;     // SLOT: [n] index of the table slot every row matched (0xffffffff: no match) -- what b200j_gather reads
                                  //       the slot-ordered build columns with (B200J_OUT_GATHER_HINT)
    uint32_t        zero_slot_index;   // SLOT: index that stands for the side bucket of key value 0
};

template <int KW, bool MARK, bool L1, bool SLOT = false>
__global__ void __launch_bounds__(PD_THREADS, B200J_SEG_MIN_CTAS) probe_dense_seg_kernel(ProbeArgs a, SegArgs sg) {
    static_assert(!(SLOT && B200J_SEG_FLAT), "the slot output is written by the segment-by-segment walk only");
    using KeyT = typename std::conditional<KW == 8, unsigned long long, uint32_t>::type;
    constexpr uint32_t FULL = 0xffffffffu;
    __shared__ PDShared sh;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const KeyT *keys = reinterpret_cast<const KeyT *>(a.keys);
    unsigned long long *out = reinterpret_cast<unsigned long long *>(a.out_build);
    unsigned long long *qkey = sh.qkey[warp];
    uint32_t *qidx = sh.qidx[warp];
    uint32_t qn = 0, misses = 0;
    const uint32_t P = 1u << a.tv.radix_bits;
    const uint32_t n_claims = (sg.n_units + SEG_UNITS - 1) / SEG_UNITS;
    uint32_t claim = 0;
    if (lane == 0) claim = atomicAdd(&a.ctr->next_tile, 1u);
    for (;;) {
        const uint32_t cur = __shfl_sync(FULL, claim, 0);
        if (cur >= n_claims) break;
        if (lane == 0) claim = atomicAdd(&a.ctr->next_tile, 1u);
        // ---- lane j: bounds and table of unit j of the claim. The claim's segments are then walked as ONE stream of `total`
        //      virtual rows: virtual row v lies in the segment s with cum[s - 1] <= v < cum[s].
        uint32_t my_start = 0, my_len = 0, my_base = 0, my_nb = 0;
        {
            const uint32_t u = cur * SEG_UNITS + (uint32_t)lane;
            if (lane < SEG_UNITS && u < sg.n_units) {
                const uint32_t p = u / sg.n_blocks, b = u - p * sg.n_blocks;
                const uint2 e = make_uint2(__ldg(sg.seg + (size_t)b * (P + 2) + p), __ldg(sg.seg + (size_t)b * (P + 2) + p + 1));
                my_start = __ldg(sg.block_base + b) + e.x;
                my_len = e.y - e.x;
                if (p < P) {
                    const uint2 pe = __ldg(reinterpret_cast<const uint2 *>(a.tv.parts + p));
                    my_base = pe.x;
                    my_nb = pe.y;
                }                                               // p == P: NULL keys, nb stays 0
            }
        }
#if B200J_SEG_FLAT
        uint32_t my_cum = my_len;                               // inclusive prefix of the lengths over the lanes
#pragma unroll
        for (int o = 1; o < SEG_UNITS; o <<= 1) { const uint32_t y = __shfl_up_sync(FULL, my_cum, o); if (lane >= o) my_cum += y; }
        const uint32_t total = __shfl_sync(FULL, my_cum, SEG_UNITS - 1);
        uint32_t cum[SEG_UNITS];                                // warp-uniform copies: the segment search is SEG_UNITS compares
#pragma unroll
        for (int j = 0; j < SEG_UNITS; j++) cum[j] = __shfl_sync(FULL, my_cum, j);
        const uint32_t my_delta = my_start - (my_cum - my_len);     // position of virtual row v of my segment = v + my_delta

        // keys of one 128-row iteration: position of every row (0xffffffff past the end) and its key, loads in flight
        auto load_iter = [&](uint32_t v0, uint32_t (&pos)[PD_ITEMS], KeyT (&k)[PD_ITEMS]) {
#pragma unroll
            for (int q = 0; q < PD_ITEMS; q++) {
                const uint32_t v = v0 + q * 32 + lane;
                uint32_t s = 0;
#pragma unroll
                for (int j = 0; j < SEG_UNITS - 1; j++) s += (v >= cum[j]) ? 1u : 0u;
                const uint32_t d = __shfl_sync(FULL, my_delta, s);
                pos[q] = v < total ? v + d : 0xffffffffu;
                k[q] = v < total ? __ldcs(keys + pos[q]) : (KeyT)0;
            }
        };
        uint32_t pos[PD_ITEMS], npos[PD_ITEMS];
        KeyT kraw[PD_ITEMS], nk[PD_ITEMS];
        load_iter(0, npos, nk);
        for (uint32_t v0 = 0; v0 < total; v0 += PD_WTILE) {
            uint64_t key[PD_ITEMS];
#pragma unroll
            for (int q = 0; q < PD_ITEMS; q++) { pos[q] = npos[q]; kraw[q] = nk[q]; }
            // the next iteration's keys are requested before this iteration's lookups are consumed (ncu, first version: 42 % of
            // all stall samples sat on the first use of a key)
            if (v0 + PD_WTILE < total) load_iter(v0 + PD_WTILE, npos, nk);
            unsigned long long k0[PD_ITEMS], v0_[PD_ITEMS], k1[PD_ITEMS], v1[PD_ITEMS];
            bool inb[PD_ITEMS], nullrow[PD_ITEMS];
#pragma unroll
            for (int q = 0; q < PD_ITEMS; q++) {
                const uint32_t v = v0 + q * 32 + lane;
                uint32_t s = 0;
#pragma unroll
                for (int j = 0; j < SEG_UNITS - 1; j++) s += (v >= cum[j]) ? 1u : 0u;
                const uint32_t base = __shfl_sync(FULL, my_base, s), nb = __shfl_sync(FULL, my_nb, s);
                inb[q] = pos[q] != 0xffffffffu;
                nullrow[q] = inb[q] && nb == 0u;
                key[q] = (uint64_t)kraw[q];
                const uint64_t h = hash_one(a.hk, key[q]);
                const Slot *p = (key[q] && nb) ? bucket_ptr(a.tv, base, home_bucket(h, nb)) : a.tv.zero_slot;
                load_bucket(p, k0[q], v0_[q], k1[q], v1[q], L1);
            }
            bool pend[PD_ITEMS];
            uint32_t n_pend = 0;
#pragma unroll
            for (int q = 0; q < PD_ITEMS; q++) {
                const bool hit0 = k0[q] == key[q], hit1 = k1[q] == key[q];
                const unsigned long long v = hit0 ? v0_[q] : v1[q];
                const bool hit = !nullrow[q] && (hit0 | hit1) && ((uint32_t)(v >> 32) & COUNT_MASK) != 0u;
                pend[q] = inb[q] && !nullrow[q] && !(hit0 | hit1) && k1[q] != 0ULL && bucket_overflowed(v0_[q]);
                n_pend += pend[q] ? 1u : 0u;
                if (inb[q] && !pend[q]) {
                    __stcs(out + pos[q], hit ? (unsigned long long)(uint32_t)v : ~0ULL);
                    if (!hit) misses++;
                    else if (MARK) mark_visited(a.visited, (uint32_t)v);
                }
            }
            const uint32_t b0 = __ballot_sync(FULL, n_pend & 1u), b1 = __ballot_sync(FULL, n_pend & 2u), b2 = __ballot_sync(FULL, n_pend & 4u);
            const uint32_t wtot = __popc(b0) + 2u * __popc(b1) + 4u * __popc(b2);
            if (wtot) {
                const uint32_t lt = (1u << lane) - 1u;
                uint32_t qp = qn + __popc(b0 & lt) + 2u * __popc(b1 & lt) + 4u * __popc(b2 & lt);
#pragma unroll
                for (int q = 0; q < PD_ITEMS; q++) {
                    if (!pend[q]) continue;
                    qkey[qp] = key[q];
                    qidx[qp] = pos[q];
                    qp++;
                }
                qn += wtot;
                __syncwarp();
                while (qn >= 32u) {
                    qn -= 32u;
                    dense_drain<MARK, true, L1>(a, qkey, qidx, qn, 32u, lane, misses);
                    __syncwarp();
                }
            }
        }
    }
#else
#pragma unroll 1
        for (int s = 0; s < SEG_UNITS; s++) {
            const uint32_t len = __shfl_sync(FULL, my_len, s);
            if (len == 0) continue;
            const uint32_t start = __shfl_sync(FULL, my_start, s);
            const uint32_t base = __shfl_sync(FULL, my_base, s), nb = __shfl_sync(FULL, my_nb, s);
            if (nb == 0) {                                      // NULL-key rows
                for (uint32_t i = lane; i < len; i += 32) {
                    __stcs(out + start + i, ~0ULL);
                    if (SLOT) __stcs(sg.out_slot + start + i, 0xffffffffu);
                }
                misses += (len + 31u - (uint32_t)lane) / 32u;
                continue;
            }
            for (uint32_t i0 = 0; i0 < len; i0 += PD_WTILE) {
                uint64_t key[PD_ITEMS];
                bool inb[PD_ITEMS];
#pragma unroll
                for (int q = 0; q < PD_ITEMS; q++) {
                    inb[q] = i0 + q * 32 + lane < len;
                    key[q] = inb[q] ? (uint64_t)__ldcs(keys + start + i0 + q * 32 + lane) : 0ULL;
                }
                unsigned long long k0[PD_ITEMS], v0[PD_ITEMS], k1[PD_ITEMS], v1[PD_ITEMS];
                uint32_t sl[SLOT ? PD_ITEMS : 1];               // SLOT: index of the even slot of the bucket looked at
#pragma unroll
                for (int q = 0; q < PD_ITEMS; q++) {
                    const uint64_t h = hash_one(a.hk, key[q]);
                    const uint32_t hb = home_bucket(h, nb);
                    const Slot *p = key[q] ? bucket_ptr(a.tv, base, hb) : a.tv.zero_slot;
                    if constexpr (SLOT) sl[q] = key[q] ? 2u * (base + hb) : sg.zero_slot_index;
                    load_bucket(p, k0[q], v0[q], k1[q], v1[q], L1);
                }
                bool pend[PD_ITEMS];
                uint32_t n_pend = 0;
#pragma unroll
                for (int q = 0; q < PD_ITEMS; q++) {
                    const bool hit0 = k0[q] == key[q], hit1 = k1[q] == key[q];
                    const unsigned long long v = hit0 ? v0[q] : v1[q];
                    const bool hit = (hit0 | hit1) && ((uint32_t)(v >> 32) & COUNT_MASK) != 0u;
                    pend[q] = inb[q] && !(hit0 | hit1) && k1[q] != 0ULL && bucket_overflowed(v0[q]);
                    n_pend += pend[q] ? 1u : 0u;
                    if (inb[q] && !pend[q]) {
                        __stcs(out + start + i0 + q * 32 + lane, hit ? (unsigned long long)(uint32_t)v : ~0ULL);
                        if constexpr (SLOT) __stcs(sg.out_slot + start + i0 + q * 32 + lane, hit ? sl[q] + (hit0 ? 0u : 1u) : 0xffffffffu);
                        if (!hit) misses++;
                        else if (MARK) mark_visited(a.visited, (uint32_t)v);
                    }
                }
                const uint32_t b0 = __ballot_sync(FULL, n_pend & 1u), b1 = __ballot_sync(FULL, n_pend & 2u), b2 = __ballot_sync(FULL, n_pend & 4u);
                const uint32_t wtot = __popc(b0) + 2u * __popc(b1) + 4u * __popc(b2);
                if (wtot) {
                    const uint32_t lt = (1u << lane) - 1u;
                    uint32_t pos = qn + __popc(b0 & lt) + 2u * __popc(b1 & lt) + 4u * __popc(b2 & lt);
#pragma unroll
                    for (int q = 0; q < PD_ITEMS; q++) {
                        if (!pend[q]) continue;
                        qkey[pos] = key[q];
                        qidx[pos] = start + i0 + q * 32 + lane;
                        pos++;
                    }
                    qn += wtot;
                    __syncwarp();
                    while (qn >= 32u) {
                        qn -= 32u;
                        dense_drain<MARK, true, L1, SLOT>(a, qkey, qidx, qn, 32u, lane, misses, sg.out_slot);
                        __syncwarp();
                    }
                }
            }
        }
    }
#endif
    if (qn) dense_drain<MARK, true, L1, SLOT>(a, qkey, qidx, 0u, qn, lane, misses, sg.out_slot);
    const unsigned long long z = warp_sum_u32(misses);
    if (lane == 0 && z) atomicAdd(&a.ctr->build_null_out, z);
}

// Compaction of a dense result: keeps the rows whose build index is not NULL. One global atomic per tile.
// Rows are read as coalesced 16 B / 8 B vectors (lane-contiguous), the survivors are ranked with one block scan, staged in
// shared memory in rank order and written out as contiguous runs, so both directions move whole sectors. The order of the
// surviving pairs inside a tile is the vector order, not the row order (the result is a multiset).
constexpr int CP_THREADS = 256;
constexpr int CP_ITEMS   = 8;
constexpr int CP_TILE    = CP_THREADS * CP_ITEMS;
constexpr int CP_VECS    = CP_ITEMS / 2;
__global__ void __launch_bounds__(CP_THREADS) compact_pairs_kernel(const uint64_t *__restrict__ in_build, const uint32_t *__restrict__ in_probe,
                                                                   uint64_t n, uint64_t *__restrict__ out_build, uint32_t *__restrict__ out_probe,
                                                                   unsigned long long *counter) {
    __shared__ unsigned long long s_b[CP_TILE];
    __shared__ uint32_t s_p[CP_TILE];
    __shared__ uint32_t s_w[CP_THREADS / 32];
    __shared__ unsigned long long s_base;
    __shared__ uint32_t s_tot;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const uint64_t n_tiles = (n + CP_TILE - 1) / CP_TILE;
    for (uint64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        unsigned long long b[CP_ITEMS];
        uint32_t p[CP_ITEMS], c = 0;
#pragma unroll
        for (int v = 0; v < CP_VECS; v++) {
            const uint64_t r0 = tile * CP_TILE + ((uint64_t)v * CP_THREADS + (uint64_t)t) * 2;
            if (r0 + 1 < n) {
                const ulonglong2 x = __ldcs(reinterpret_cast<const ulonglong2 *>(in_build + r0));
                const uint2 y = __ldcs(reinterpret_cast<const uint2 *>(in_probe + r0));
                b[2 * v] = x.x; b[2 * v + 1] = x.y;
                p[2 * v] = y.x; p[2 * v + 1] = y.y;
            } else {
                const bool in = r0 < n;
                b[2 * v] = in ? __ldcs(reinterpret_cast<const unsigned long long *>(in_build) + r0) : ~0ULL;
                p[2 * v] = in ? __ldcs(in_probe + r0) : 0u;
                b[2 * v + 1] = ~0ULL;
                p[2 * v + 1] = 0u;
            }
            c += (b[2 * v] != ~0ULL ? 1u : 0u) + (b[2 * v + 1] != ~0ULL ? 1u : 0u);
        }
        uint32_t incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
        if (lane == 31) s_w[warp] = incl;
        __syncthreads();
        if (t == 0) {
            uint32_t tot = 0;
            for (int i = 0; i < CP_THREADS / 32; i++) tot += s_w[i];
            s_tot = tot;
            s_base = tot ? atomicAdd(counter, (unsigned long long)tot) : 0ULL;
        }
        uint32_t pos = incl - c;
        for (int i = 0; i < warp; i++) pos += s_w[i];
#pragma unroll
        for (int q = 0; q < CP_ITEMS; q++) {
            if (b[q] == ~0ULL) continue;
            s_b[pos] = b[q];
            s_p[pos] = p[q];
            pos++;
        }
        __syncthreads();
        const uint32_t tot = s_tot;
        const unsigned long long base = s_base;
        for (uint32_t i = (uint32_t)t; i < tot; i += CP_THREADS) {
            __stcs(reinterpret_cast<unsigned long long *>(out_build) + base + i, s_b[i]);
            __stcs(out_probe + base + i, s_p[i]);
        }
        __syncthreads();
    }
}

}  // namespace b200j
