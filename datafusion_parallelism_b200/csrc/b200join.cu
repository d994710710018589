// b200join.cu — host side of libb200join.so: the C ABI of include/b200join.h over the sm_100a kernels in
// kernels.cuh / radix.cuh / gather.cuh.  C++17, CUDA runtime only (no torch, no arrow, no CPU fallback).
//
// Mirrors, for one join, what sits below ParallelHashJoin::execute in the reference
// (src/operator/parallel_hash_join.rs:140-167 -> parallel_hash_join_executor.rs:86-122):
// build_side() per partition, then one lookup_*_probe_batch per probe batch, then the finaliser.
#include "../../include/b200join.h"
#include "kernels.cuh"
#include "radix.cuh"
#include "radix_block.cuh"
#include "probe.cuh"
#include "gather.cuh"
#include "gather_remote.cuh"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <ctime>
#include <mutex>
#include <set>
#include <string>
#include <vector>

using namespace b200j;

namespace {

thread_local std::string g_last_error;

struct Fail {
    int32_t code;
    std::string msg;
};

#define CU(expr)                                                                                          \
    do {                                                                                                  \
        cudaError_t _e = (expr);                                                                          \
        if (_e != cudaSuccess) {                                                                          \
            cudaGetLastError();                                                                           \
            throw Fail{_e == cudaErrorMemoryAllocation ? B200J_ERR_OOM : B200J_ERR_CUDA,                  \
                       std::string(#expr) + ": " + cudaGetErrorString(_e)};                               \
        }                                                                                                 \
    } while (0)

#define REQUIRE(cond, code, text)                                                                         \
    do {                                                                                                  \
        if (!(cond)) throw Fail{(code), std::string(text)};                                               \
    } while (0)

// RandomState::from_keys over the fixed PI seeds (ahash random_state.rs), identical to oracle/join_oracle.c.
void keys_with_seed(uint64_t seed, uint64_t out[4]) {
    static const uint64_t PI[4] = {0x243f6a8885a308d3ULL, 0x13198a2e03707344ULL, 0xa4093822299f31d0ULL, 0x082efa98ec4e6c89ULL};
    static const uint64_t PI2[4] = {0x452821e638d01377ULL, 0xbe5466cf34e90c6cULL, 0xc0ac29b7c97c50ddULL, 0x3f84d5b5b5470917ULL};
    uint64_t buffer = PI[1];
    const uint64_t pad = PI[0];
    buffer = folded_multiply(seed ^ buffer, AHASH_MULTIPLE);
    const int l[4] = {0, 1, 2, 3}, r[4] = {2, 3, 1, 0};
    for (int i = 0; i < 4; i++) {
        uint64_t b = buffer;
        b = folded_multiply(PI2[l[i]] ^ b, AHASH_MULTIPLE);
        b = folded_multiply(PI2[r[i]] ^ b, AHASH_MULTIPLE);
        out[i] = rotl64(folded_multiply(b, pad), (unsigned)(b & 63));
    }
}

struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
};

struct ColStore {          // one column of the concatenated build side (JoinStateInstance::compact, a3)
    int32_t type = 8;      // width or B200J_T_UTF8
    DevBuf values;         // fixed width values, or Utf8 data
    DevBuf offsets;        // Utf8: int64 offsets into `values` (the concatenated build side may pass 2 GiB)
    DevBuf validity;       // u32 words, allocated on the first batch that carries nulls
    bool has_nulls = false;
    int64_t data_bytes = 0;
};

// b200j_result grew at its end (build_idx_width): callers built against the first layout stay valid
constexpr size_t RESULT_V1_SIZE = offsetof(b200j_result, build_idx_width);

struct ResultOwner {
    void *build_idx = nullptr;
    void *probe_idx = nullptr;
    int32_t mem = B200J_MEM_DEVICE;
    int32_t bwidth = 8;               // bytes per build index: 4 after B200J_OUT_BUILD_U32 narrowing
    bool probe_order = false;         // B200J_OUT_PROBE_ORDER honoured: build_idx[i] belongs to probe row i, probe_idx is null
    // The dense probe behind the radix layout leaves the pairs in the order of the block-partitioned probe stream. When the
    // caller of probe_device asks for it (want_seg), the stream's segment table stays with the result so that b200j_gather can
    // walk the build side partition by partition (gather_seg_multi_kernel). Device memory, whatever `mem` says.
    bool want_seg = false;
    void *seg = nullptr;              // [seg_blocks][seg_P + 2] segment starts, then [seg_blocks] block bases
    uint32_t seg_blocks = 0, seg_P = 0;
    int64_t seg_n = 0;                // rows the table describes (= the result's rows)
    // B200J_OUT_GATHER_HINT (want_aux): what else the dense probe can leave behind for the gather
    bool want_aux = false;
    void *aux_keys = nullptr;         // the partitioned probe keys = the probe key of every result row, in result order (no NULL keys in the batch)
    int32_t aux_kw = 0;               // ... 4 or 8 bytes each (4: zero-extended values of 8-byte keys)
    void *aux_slot = nullptr;         // u32 per result row: index of the table slot the row matched (0xffffffff: none)
    int64_t aux_misses = 0;           // result rows without a match (NULL build index)
};

struct CachedBlock {
    void *p = nullptr;
    size_t cap = 0;
    bool in_use = false;
};

struct PartPlan {            // one planned partitioning pass (radix_host.inl)
    PartArgs a{};
    int grid = 1;
    int kw = 8;
    bool dest = false;
};

struct PinnedBuf {
    void *p = nullptr;
    size_t cap = 0;
};

}  // namespace

struct b200j_handle {
    std::mutex mu;
    std::string err;
    b200j_config cfg{};
    std::vector<int32_t> key_widths, build_payload_types, probe_payload_types;
    HashKeys hk{};
    cudaStream_t stream = nullptr;
    cudaStream_t copy_in = nullptr, copy_out = nullptr;
    bool own_stream = false;
    int sm_count = 148;
    int probe_ctas_per_sm = 4;
    int dense_ctas_per_sm = 4;
    int seg_ctas_per_sm = 4;
    bool fused_build = false;          // radix layout: table clear + insert in one persistent kernel (build_insert_fused_kernel). Measured NEUTRAL at C2
                                       // (insert phase 3.35 vs 3.41 ms: the insert is not waiting on DRAM), so the memset + build_insert_kernel
                                       // path without a spin-wait stays the default; B200J_FUSED_BUILD=1 selects it
    int fused_ctas_per_sm = 5;
    bool narrow_ok = true;             // 8-byte probe keys go through the partitioned stream as 32-bit values until one does not fit (B200J_NO_NARROW=1: never)
    bool block_part_ok = true;         // B200J_NO_BLOCK_PART=1: the dense radix probe keeps the two-pass global partition of round 1
    int pb_clusters[2][PB_MAX_BITS + 1] = {};   // resident clusters of part_block_kernel per key width / radix bits (0 = not asked yet)
    unsigned long long *d_part_rows = nullptr, *h_part_rows = nullptr;   // rows per partition of the last block-partitioned batch
    uint32_t *d_reserve = nullptr;     // region cursors of b200j_shuffle_local_scatter
    cudaStream_t lanes[4] = {nullptr, nullptr, nullptr, nullptr};        // copy lanes of b200j_copy_async
    bool dense_ok = true;              // B200J_NO_DENSE=1 forces the staged probe kernel (tuning / tests)
    uint64_t l1_table_bytes = 512 << 10; // direct tables up to this size are probed through L1 (B200J_L1_TABLE_BYTES)
    float thin_lf_min = 0.2f;          // lowest load factor the auto-thinning of small direct tables goes to (B200J_THIN_LF_MIN; 0.5 = off)
    uint64_t thin_table_bytes = 48ull << 20;   // ... while the table stays within this many bytes (B200J_THIN_TABLE_BYTES)
    bool skew_l1 = true;               // radix dense probe: L1-allocating bucket loads for skewed batches (B200J_SKEW_L1=0 turns it off)
    float radix_lf = 0.42f;            // load factor of the radix layout's tables when the caller leaves it open (B200J_RADIX_LF)
    float miss_frac = 0.f;             // share of probe rows without a match in the last pair-producing probe launch
    float dense_miss_max = 0.3f;       // direct layout: above this miss share the staged kernel is used (B200J_DENSE_MISS_MAX)
    int64_t l2_bytes = 0;
    bool timing = false;

    // build store
    int64_t n_build = 0;
    int64_t build_cap_rows = 0;
    std::vector<ColStore> bkeys, bpayload;
    std::vector<uint8_t> build_done, probe_done;
    int n_build_done = 0, n_probe_done = 0;
    bool built = false;
    bool composite = false;            // n_keys > 1: the table is keyed on the 64-bit row hash, equality checked on candidate pairs
    int table_kw = 8;                  // width of what the table stores: the key column's, or 8 (row hash)

    // coalescing feeder: one stage per partition (b200j_probe_push)
    struct ProbeStage {
        std::vector<ColStore> keys, payload;
        int64_t n = 0, cap_rows = 0;
    };
    std::vector<ProbeStage> stages;

    // table
    int32_t layout = B200J_LAYOUT_DIRECT;
    int32_t radix_bits = 0;
    DevBuf slots;
    uint64_t n_slots = 0;
    DevBuf zero_slot;
    DevBuf rowids;
    DevBuf hot_table;                  // heavy-hitter key set of the exchange (b200j_shuffle_set_hot)
    uint32_t hot_mask = 0;
    int hot_n = 0;
    DevBuf bigruns;                    // BigRun list of the probe launch in flight (allocated when the build has long duplicate runs)
    DevBuf visited;
    DevBuf parts;                      // radix layout: PartEnt per partition
    RadixScratch rscratch;
    uint32_t probe_rid_base = 0;       // chunked host probe: row id of the chunk's first row inside the caller's batch
    bool shuffle_armed = false;        // b200j_shuffle_count done, b200j_shuffle_scatter pending
    PartPlan shuffle_plan;
    int64_t shuffle_n = 0;
    // slot-ordered copies of the fixed-width build payload columns (B200J_OUT_GATHER_HINT): value of the row held by table slot s at
    // index s, [n_slots + 2] entries (the last two: the side bucket of key 0); built on the first hinted gather after a build
    struct SlotCol { DevBuf values, validity; bool has_nulls = false; };
    std::vector<SlotCol> slot_cols;
    bool slot_cols_ready = false;
    int64_t probed_since_build = 0;    // probe rows that went through the dense radix path since the table was built
    int utf8_lanes = 0;                // lanes per row of the Utf8 gather copy (0 = from the average value length; B200J_UTF8_LANES)
    bool auto_hint = true;             // a handle whose device results have been gathered once treats later device probes as hinted (B200J_AUTO_HINT=0: off)
    bool gather_seen = false;
    bool gather_seg = true;            // b200j_gather: build columns of a dense radix result are gathered partition by partition (B200J_GATHER_SEG=0: off)
    std::set<const void *> seg_owners;  // live results that carry a segment table (b200j_gather checks a caller's _owner against it)
    bool key_via_probe = true;         // b200j_gather: the build key column of matched pairs is read from the probe key column (B200J_KEY_VIA_PROBE=0: off)
    uint32_t *static_hist = nullptr;   // device address of g_static_hist ON THIS HANDLE'S DEVICE (b200j_shuffle_count_static)
    Counters *d_ctr = nullptr;
    Counters *h_ctr = nullptr;         // pinned mirror

    std::vector<PinnedBuf> pinned_cache;
    std::vector<CachedBlock> blocks;   // device buffers >= 1 MiB, reused across calls (see big_alloc)
    size_t blocks_free_bytes = 0;
    size_t cache_limit = (size_t)64 << 30;   // free cached bytes kept at most (set in b200j_create)
    bool guard = false;                // B200J_GUARD=1: every block of >= 1 MiB sits between two 256-byte canaries that are checked when it is
                                       // freed (compute-sanitizer is closed on this GPU pool: this is the library's own out-of-bounds check)
    long guard_violations = 0;
    b200j_stats st{};
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;

    TableView view() const {
        TableView tv{};
        tv.slots = (Slot *)slots.p;
        tv.zero_slot = (Slot *)zero_slot.p;
        tv.rowids = (uint32_t *)rowids.p;
        tv.parts = (const PartEnt *)parts.p;
        tv.nb = (uint32_t)(n_slots / 2);
        tv.radix_bits = radix_bits;
        tv.l1 = (radix_bits == 0 && n_slots * 16 <= l1_table_bytes) ? 1u : 0u;
        return tv;
    }
};

namespace {

using H = b200j_handle;

// defined in radix_host.inl / gather_host.inl (included below, they need the complete handle type)
void radix_build(H *h, const void *keys, const uint8_t *validity, int kw, int64_t n, float lf, int64_t budget);
void radix_partition(H *h, const void *keys, const uint8_t *validity, int64_t bit_offset, int kw, int64_t n, int bits,
                     void *out_keys, uint32_t *out_rid, double *ms_acc);
void radix_release(H *h);
struct ProbePlan;
struct ProbePiece;
int64_t probe_dense_blocked_pieces(H *h, const ProbePiece *pieces, int n_pieces, const ProbePlan &pl, ResultOwner *ro, int64_t *n_build_null,
                                   int piece_kw = 0);
int64_t probe_dense_blocked(H *h, const void *d_keys, const uint8_t *d_validity, int64_t bit_offset, int64_t n, const ProbePlan &pl,
                            ResultOwner *ro, int64_t *n_build_null);
void do_gather(H *h, const b200j_result *pairs, const b200j_batch *probe_batch, uint32_t want, b200j_gathered *out);
void do_gathered_release(H *h, b200j_gathered *g);

// Device allocation for everything a join touches. Small blocks go to the stream-ordered pool. Large ones are cached in
// the handle: measured on B200, cudaMallocAsync of an 8 GB block cost 6.7 ms per probe call even with the pool's
// release threshold at its maximum, more than the probe kernel itself. All work of a handle is ordered on one stream, so
// a block freed here may be handed out again at once.
constexpr size_t BIG_BLOCK = (size_t)1 << 20;
// Free cached blocks are kept up to h->cache_limit bytes (40 % of the device's memory unless B200J_CACHE_LIMIT_BYTES says
// otherwise); b200j_trim gives all of them back at once (another allocator in the process may need the room).

constexpr size_t GUARD_BYTES = 256;
constexpr int GUARD_PATTERN = 0xA5;
void guard_check(H *h, void *user_ptr, size_t user_bytes);

template <typename T>
cudaError_t big_alloc_raw(H *h, T **out, size_t bytes);
cudaError_t big_free_raw(H *h, void *p);

struct GuardRec { void *user; size_t bytes; };
std::mutex g_guard_mu;
std::vector<GuardRec> g_guard_recs;

template <typename T>
cudaError_t big_alloc(H *h, T **out, size_t bytes) {
    if (!h->guard) return big_alloc_raw(h, out, bytes);
    *out = nullptr;
    const size_t ub = (bytes + 15) & ~(size_t)15;
    char *raw = nullptr;
    cudaError_t e = big_alloc_raw(h, &raw, ub + 2 * GUARD_BYTES);
    if (e != cudaSuccess) return e;
    cudaMemsetAsync(raw, GUARD_PATTERN, GUARD_BYTES, h->stream);
    cudaMemsetAsync(raw + GUARD_BYTES + ub, GUARD_PATTERN, GUARD_BYTES, h->stream);
    *out = (T *)(raw + GUARD_BYTES);
    std::lock_guard<std::mutex> g(g_guard_mu);
    g_guard_recs.push_back(GuardRec{raw + GUARD_BYTES, ub});
    return cudaSuccess;
}

cudaError_t big_free(H *h, void *p) {
    if (!p || !h->guard) return big_free_raw(h, p);
    size_t ub = 0;
    bool found = false;
    {
        std::lock_guard<std::mutex> g(g_guard_mu);
        for (size_t i = 0; i < g_guard_recs.size(); i++)
            if (g_guard_recs[i].user == p) { ub = g_guard_recs[i].bytes; g_guard_recs.erase(g_guard_recs.begin() + i); found = true; break; }
    }
    if (!found) return big_free_raw(h, p);          // allocated before the guard was switched on
    guard_check(h, p, ub);
    return big_free_raw(h, (char *)p - GUARD_BYTES);
}

template <typename T>
cudaError_t big_alloc_raw(H *h, T **out, size_t bytes) {
    *out = nullptr;
    if (bytes < BIG_BLOCK) return cudaMallocAsync((void **)out, bytes, h->stream);
    int best = -1;
    for (size_t i = 0; i < h->blocks.size(); i++) {
        const CachedBlock &b = h->blocks[i];
        if (!b.in_use && b.cap >= bytes && b.cap <= bytes + bytes / 4 && (best < 0 || b.cap < h->blocks[best].cap)) best = (int)i;
    }
    if (best >= 0) {
        h->blocks[best].in_use = true;
        h->blocks_free_bytes -= h->blocks[best].cap;
        *out = (T *)h->blocks[best].p;
        return cudaSuccess;
    }
    void *p = nullptr;
    cudaError_t e = cudaMallocAsync(&p, bytes, h->stream);
    if (e != cudaSuccess) {
        // make room: drop every cached free block and try once more
        cudaGetLastError();
        for (size_t i = 0; i < h->blocks.size();) {
            if (!h->blocks[i].in_use) {
                cudaFreeAsync(h->blocks[i].p, h->stream);
                h->blocks_free_bytes -= h->blocks[i].cap;
                h->blocks.erase(h->blocks.begin() + i);
            } else i++;
        }
        cudaStreamSynchronize(h->stream);
        e = cudaMallocAsync(&p, bytes, h->stream);
        if (e != cudaSuccess) return e;
    }
    h->blocks.push_back(CachedBlock{p, bytes, true});
    *out = (T *)p;
    return cudaSuccess;
}

cudaError_t big_free_raw(H *h, void *p) {
    if (!p) return cudaSuccess;
    for (size_t i = 0; i < h->blocks.size(); i++)
        if (h->blocks[i].p == p) {
            h->blocks[i].in_use = false;
            h->blocks_free_bytes += h->blocks[i].cap;
            // bounded cache: give the largest free blocks back to the pool beyond the limit
            while (h->blocks_free_bytes > h->cache_limit) {
                int big = -1;
                for (size_t j = 0; j < h->blocks.size(); j++)
                    if (!h->blocks[j].in_use && (big < 0 || h->blocks[j].cap > h->blocks[big].cap)) big = (int)j;
                if (big < 0) break;
                cudaFreeAsync(h->blocks[big].p, h->stream);
                h->blocks_free_bytes -= h->blocks[big].cap;
                h->blocks.erase(h->blocks.begin() + big);
            }
            return cudaSuccess;
        }
    return cudaFreeAsync(p, h->stream);
}

void guard_check(H *h, void *user_ptr, size_t user_bytes) {
    unsigned char host[2 * GUARD_BYTES];
    cudaStreamSynchronize(h->stream);
    if (cudaMemcpy(host, (char *)user_ptr - GUARD_BYTES, GUARD_BYTES, cudaMemcpyDeviceToHost) != cudaSuccess ||
        cudaMemcpy(host + GUARD_BYTES, (char *)user_ptr + user_bytes, GUARD_BYTES, cudaMemcpyDeviceToHost) != cudaSuccess) {
        cudaGetLastError();
        return;
    }
    for (size_t i = 0; i < 2 * GUARD_BYTES; i++)
        if (host[i] != GUARD_PATTERN) {
            h->guard_violations++;
            fprintf(stderr, "[b200j] GUARD VIOLATION: %s canary of a %zu-byte block overwritten at byte %zu\n", i < GUARD_BYTES ? "front" : "back",
                    user_bytes, i % GUARD_BYTES);
            break;
        }
}

// Gives every cached free block back to the stream-ordered pool and trims the pool itself (b200j_trim).
void cache_trim(H *h) {
    for (size_t i = 0; i < h->blocks.size();) {
        if (!h->blocks[i].in_use) {
            cudaFreeAsync(h->blocks[i].p, h->stream);
            h->blocks_free_bytes -= h->blocks[i].cap;
            h->blocks.erase(h->blocks.begin() + i);
        } else i++;
    }
    CU(cudaStreamSynchronize(h->stream));
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, h->cfg.device) == cudaSuccess) cudaMemPoolTrimTo(pool, 0);
    cudaGetLastError();
}

void dev_alloc(H *h, DevBuf &b, size_t bytes) {
    if (bytes <= b.cap && b.p) return;
    if (b.p) CU(big_free(h, b.p));
    b.p = nullptr;
    b.cap = 0;
    bytes = std::max<size_t>(bytes, 256);
    CU(big_alloc(h, &b.p, bytes));
    b.cap = bytes;
}

// grow keeping the first `keep` bytes
void dev_grow(H *h, DevBuf &b, size_t bytes, size_t keep) {
    if (bytes <= b.cap && b.p) return;
    size_t ncap = std::max<size_t>(std::max<size_t>(bytes, b.cap + b.cap / 2), 256);
    void *np = nullptr;
    CU(big_alloc(h, &np, ncap));
    if (b.p && keep) CU(cudaMemcpyAsync(np, b.p, keep, cudaMemcpyDeviceToDevice, h->stream));
    if (b.p) CU(big_free(h, b.p));
    b.p = np;
    b.cap = ncap;
}

void dev_free(H *h, DevBuf &b) {
    if (b.p) big_free(h, b.p);
    b.p = nullptr;
    b.cap = 0;
}

struct PinnedSizes {
    std::mutex mu;
    std::vector<std::pair<void *, size_t>> v;
} g_pinned_sizes;

void pinned_note(void *p, size_t bytes) {
    std::lock_guard<std::mutex> g(g_pinned_sizes.mu);
    g_pinned_sizes.v.emplace_back(p, bytes);
}
size_t pinned_size(void *p, bool erase) {
    std::lock_guard<std::mutex> g(g_pinned_sizes.mu);
    for (size_t i = 0; i < g_pinned_sizes.v.size(); i++)
        if (g_pinned_sizes.v[i].first == p) {
            size_t s = g_pinned_sizes.v[i].second;
            if (erase) g_pinned_sizes.v.erase(g_pinned_sizes.v.begin() + i);
            return s;
        }
    return 0;
}

void *pinned_take(H *h, size_t bytes) {
    bytes = std::max<size_t>(bytes, 256);
    int best = -1;
    for (size_t i = 0; i < h->pinned_cache.size(); i++)
        if (h->pinned_cache[i].cap >= bytes && (best < 0 || h->pinned_cache[i].cap < h->pinned_cache[best].cap)) best = (int)i;
    if (best >= 0) {
        void *p = h->pinned_cache[best].p;
        h->pinned_cache.erase(h->pinned_cache.begin() + best);
        return p;
    }
    void *p = nullptr;
    CU(cudaHostAlloc(&p, bytes, cudaHostAllocDefault));
    pinned_note(p, bytes);
    return p;
}

void pinned_give(H *h, void *p) {
    if (!p) return;
    size_t s = pinned_size(p, false);
    // keep at most 4 cached buffers; free the smallest beyond that
    h->pinned_cache.push_back(PinnedBuf{p, s});
    while (h->pinned_cache.size() > 4) {
        size_t mi = 0;
        for (size_t i = 1; i < h->pinned_cache.size(); i++)
            if (h->pinned_cache[i].cap < h->pinned_cache[mi].cap) mi = i;
        pinned_size(h->pinned_cache[mi].p, true);
        cudaFreeHost(h->pinned_cache[mi].p);
        h->pinned_cache.erase(h->pinned_cache.begin() + mi);
    }
}

int grid_for(H *h, int64_t work_items, int per_block, int blocks_per_sm) {
    int64_t blocks = (work_items + per_block - 1) / per_block;
    int64_t cap = (int64_t)h->sm_count * blocks_per_sm;
    return (int)std::max<int64_t>(1, std::min<int64_t>(blocks, cap));
}

struct Trace {                 // B200J_TRACE=1: wall-clock trace of the host side of a call (stderr)
    bool on;
    double t0;
    static double now() { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6; }
    Trace() : on(getenv("B200J_TRACE") != nullptr), t0(now()) {}
    void mark(const char *what) { if (on) { const double t = now(); fprintf(stderr, "[b200j] %-28s +%.3f ms\n", what, t - t0); t0 = t; } }
};

struct PhaseTimer {            // CUDA-event timing of one phase on the handle's stream
    H *h;
    double *acc;
    bool on;
    PhaseTimer(H *h_, double *acc_) : h(h_), acc(acc_), on(h_->timing) {
        if (on) cudaEventRecord(h->ev0, h->stream);
    }
    void stop() {
        if (!on) return;
        on = false;
        cudaEventRecord(h->ev1, h->stream);
        cudaEventSynchronize(h->ev1);
        float ms = 0;
        cudaEventElapsedTime(&ms, h->ev0, h->ev1);
        *acc += ms;
    }
};

void check_batch(H *h, const b200j_batch *b, int n_payload_expected) {
    REQUIRE(b != nullptr, B200J_ERR_INVALID, "batch is NULL");
    REQUIRE(b->struct_size >= sizeof(b200j_batch), B200J_ERR_INVALID, "batch.struct_size too small (ABI mismatch)");
    REQUIRE(b->n_rows >= 0, B200J_ERR_INVALID, "batch.n_rows < 0");
    REQUIRE(b->n_keys == h->cfg.n_keys, B200J_ERR_INVALID, "batch.n_keys differs from the join's key count");
    REQUIRE(b->n_payload == 0 || b->n_payload == n_payload_expected, B200J_ERR_INVALID,
            "batch.n_payload differs from the configured payload columns");
    REQUIRE(b->mem == B200J_MEM_HOST || b->mem == B200J_MEM_DEVICE, B200J_ERR_INVALID, "batch.mem invalid");
    if (b->n_rows > 0) {
        REQUIRE(b->keys != nullptr, B200J_ERR_INVALID, "batch.keys is NULL");
        for (int i = 0; i < b->n_keys; i++)       // a Utf8 column of empty / NULL strings only may come without a data buffer
            REQUIRE(b->keys[i].values != nullptr || h->key_widths[i] == B200J_T_UTF8, B200J_ERR_INVALID, "key column without values buffer");
    }
}

// Brings a (possibly host) validity bitmap slice to the device. Returns a device pointer whose bit 0..7
// alignment equals bit_offset & 7 (the returned *bit_off is the offset to use with it), or null.
const uint8_t *stage_validity(H *h, const b200j_column &c, int64_t n_rows, int mem, std::vector<void *> &temps, int64_t *bit_off) {
    *bit_off = 0;
    if (!c.validity || n_rows == 0) return nullptr;
    if (mem == B200J_MEM_DEVICE) {
        *bit_off = c.bit_offset;
        return c.validity;
    }
    const int64_t first = c.bit_offset >> 3, last = (c.bit_offset + n_rows + 7) >> 3;
    void *d = nullptr;
    CU(big_alloc(h, &d, (size_t)(last - first) + 16));
    temps.push_back(d);
    CU(cudaMemcpyAsync(d, c.validity + first, (size_t)(last - first), cudaMemcpyHostToDevice, h->stream));
    *bit_off = c.bit_offset & 7;
    return (const uint8_t *)d;
}

void free_temps(H *h, std::vector<void *> &temps) {
    for (void *p : temps) big_free(h, p);
    temps.clear();
}

// The key columns of a batch as the kernels' KeyCols descriptor; host columns are copied to the device (temps).
// Utf8 keys (width 0): 32-bit Arrow offsets + data bytes; only the bytes the batch's rows refer to are copied.
KeyCols stage_key_cols(H *h, const b200j_batch *b, int64_t n, std::vector<void *> &temps) {
    KeyCols kc{};
    kc.n = h->cfg.n_keys;
    for (int i = 0; i < kc.n && n > 0; i++) {
        const int w = h->key_widths[i];
        const b200j_column &c = b->keys[i];
        kc.c[i].width = w;
        kc.c[i].off_width = 4;
        if (w == B200J_T_UTF8) REQUIRE(c.offsets != nullptr, B200J_ERR_INVALID, "Utf8 key column without offsets buffer");
        if (b->mem == B200J_MEM_DEVICE) {
            kc.c[i].values = c.values;
            kc.c[i].offsets = c.offsets;
            kc.c[i].validity = c.validity;
            kc.c[i].bit_offset = c.bit_offset;
            continue;
        }
        void *d = nullptr;
        if (w == B200J_T_UTF8) {
            const int64_t o0 = c.offsets[0], nbytes = (int64_t)c.offsets[n] - o0;
            void *doff = nullptr;
            CU(big_alloc(h, &doff, (size_t)(n + 1) * 4 + 16));
            temps.push_back(doff);
            CU(cudaMemcpyAsync(doff, c.offsets, (size_t)(n + 1) * 4, cudaMemcpyHostToDevice, h->stream));
            CU(big_alloc(h, &d, (size_t)nbytes + 16));
            temps.push_back(d);
            if (nbytes) CU(cudaMemcpyAsync(d, (const char *)c.values + o0, (size_t)nbytes, cudaMemcpyHostToDevice, h->stream));
            kc.c[i].values = (const char *)d - o0;          // the offsets keep counting from the host buffer's start
            kc.c[i].offsets = doff;
        } else {
            CU(big_alloc(h, &d, (size_t)n * w + 16));
            temps.push_back(d);
            CU(cudaMemcpyAsync(d, c.values, (size_t)n * w, cudaMemcpyHostToDevice, h->stream));
            kc.c[i].values = d;
        }
        int64_t boff = 0;
        kc.c[i].validity = stage_validity(h, c, n, B200J_MEM_HOST, temps, &boff);
        kc.c[i].bit_offset = boff;
    }
    return kc;
}

// ---------------------------------------------------------------------------------------------
// build
// ---------------------------------------------------------------------------------------------
void store_append(H *h, ColStore &cs, const b200j_column &c, int64_t n_old, int64_t n_add, int mem, std::vector<void *> &temps) {
    const cudaMemcpyKind kind = mem == B200J_MEM_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice;
    if (cs.type != B200J_T_UTF8) {
        const size_t w = (size_t)cs.type;
        dev_grow(h, cs.values, (size_t)(n_old + n_add) * w + 16, (size_t)n_old * w);
        CU(cudaMemcpyAsync((char *)cs.values.p + (size_t)n_old * w, c.values, (size_t)n_add * w, kind, h->stream));
    } else {
        REQUIRE(c.offsets != nullptr, B200J_ERR_INVALID, "Utf8 column without offsets buffer");
        // the first/last offsets are needed on the host to size the copy
        int32_t o0 = 0, o1 = 0;
        if (mem == B200J_MEM_HOST) {
            o0 = c.offsets[0];
            o1 = c.offsets[n_add];
        } else {
            CU(cudaMemcpyAsync(&o0, c.offsets, 4, cudaMemcpyDeviceToHost, h->stream));
            CU(cudaMemcpyAsync(&o1, c.offsets + n_add, 4, cudaMemcpyDeviceToHost, h->stream));
            CU(cudaStreamSynchronize(h->stream));
        }
        const int64_t nbytes = (int64_t)o1 - o0;
        dev_grow(h, cs.values, (size_t)(cs.data_bytes + nbytes) + 16, (size_t)cs.data_bytes);
        if (nbytes) CU(cudaMemcpyAsync((char *)cs.values.p + cs.data_bytes, (const char *)c.values + o0, (size_t)nbytes, kind, h->stream));
        dev_grow(h, cs.offsets, (size_t)(n_old + n_add + 1) * 8, (size_t)(n_old + 1) * 8);
        const int32_t *d_off = c.offsets;
        if (mem == B200J_MEM_HOST) {
            void *d = nullptr;
            CU(big_alloc(h, &d, (size_t)(n_add + 1) * 4));
            temps.push_back(d);
            CU(cudaMemcpyAsync(d, c.offsets, (size_t)(n_add + 1) * 4, cudaMemcpyHostToDevice, h->stream));
            d_off = (const int32_t *)d;
        }
        rebase_offsets_kernel<<<grid_for(h, n_add + 1, 256, 8), 256, 0, h->stream>>>(d_off, n_add + 1, (int64_t)o0, cs.data_bytes,
                                                                                     (int64_t *)cs.offsets.p + n_old);
        h->st.kernel_launches++;
        cs.data_bytes += nbytes;
    }
    // validity
    if (c.validity && !cs.has_nulls) {
        // first batch with a bitmap: earlier rows are all valid
        dev_grow(h, cs.validity, (size_t)((std::max(h->build_cap_rows, n_old + n_add) + 31) / 32 + 1) * 4, 0);
        if (n_old > 0) {
            copy_bits_kernel<<<grid_for(h, (n_old + 31) / 32 + 1, 256, 8), 256, 0, h->stream>>>((uint32_t *)cs.validity.p, 0, nullptr, 0, n_old);
            h->st.kernel_launches++;
        }
        cs.has_nulls = true;
    }
    if (cs.has_nulls) {
        dev_grow(h, cs.validity, (size_t)((n_old + n_add + 31) / 32 + 1) * 4, (size_t)((n_old + 31) / 32) * 4);
        int64_t boff = 0;
        const uint8_t *dv = stage_validity(h, c, n_add, mem, temps, &boff);
        copy_bits_kernel<<<grid_for(h, (n_add + 31) / 32 + 1, 256, 8), 256, 0, h->stream>>>((uint32_t *)cs.validity.p, n_old, dv, boff, n_add);
        h->st.kernel_launches++;
    }
}

void do_build_push(H *h, int32_t partition, const b200j_batch *b) {
    REQUIRE(partition >= 0 && partition < h->cfg.n_partitions, B200J_ERR_INVALID, "partition out of range");
    REQUIRE(!h->build_done[partition], B200J_ERR_STATE, "State already consumed: build_push after build_finish");
    REQUIRE(!h->built, B200J_ERR_STATE, "build side already finished");
    check_batch(h, b, h->cfg.n_build_payload);
    if (b->n_rows == 0) return;
    REQUIRE(b->n_payload == h->cfg.n_build_payload, B200J_ERR_INVALID, "build batch must carry every configured payload column");
    REQUIRE(h->n_build + b->n_rows < 0x7fffffffLL, B200J_ERR_UNSUPPORTED, "more than 2^31-1 build rows on one device");
    PhaseTimer tm(h, &h->st.ms_build_copy);
    std::vector<void *> temps;
    h->build_cap_rows = std::max(h->build_cap_rows, h->n_build + b->n_rows);
    for (int i = 0; i < h->cfg.n_keys; i++) store_append(h, h->bkeys[i], b->keys[i], h->n_build, b->n_rows, b->mem, temps);
    for (int i = 0; i < h->cfg.n_build_payload; i++) store_append(h, h->bpayload[i], b->payload[i], h->n_build, b->n_rows, b->mem, temps);
    free_temps(h, temps);
    h->n_build += b->n_rows;
    // input buffers are borrowed for the call only (include/b200join.h): the copies out of them -- host OR device -- have to be
    // complete when the call returns, or a caller that frees / reuses a device batch right away corrupts the build side
    CU(cudaStreamSynchronize(h->stream));
    tm.stop();
}

template <typename F>
void launch_by_kw(int kw, F &&f) {
    if (kw == 8) f(std::integral_constant<int, 8>{});
    else f(std::integral_constant<int, 4>{});
}

// slots of a table for `n` keys at load factor lf: whole buckets, at least 8 buckets, NOT rounded to a power of two
uint64_t table_slots_for(int64_t n, float lf) {
    const uint64_t buckets = (uint64_t)std::ceil((double)n / (2.0 * (double)lf)) + 8;
    return buckets * 2;
}

// the build store's key columns as the kernels' KeyCols descriptor (row i of the store = bit i of the bitmap)
KeyCols build_key_cols(H *h) {
    KeyCols kc{};
    kc.n = h->cfg.n_keys;
    for (int i = 0; i < kc.n; i++) {
        kc.c[i].values = h->bkeys[i].values.p;
        kc.c[i].validity = h->bkeys[i].has_nulls ? (const uint8_t *)h->bkeys[i].validity.p : nullptr;
        kc.c[i].bit_offset = 0;
        kc.c[i].width = h->key_widths[i];
        kc.c[i].offsets = h->bkeys[i].offsets.p;           // Utf8: the store's 64-bit offsets
        kc.c[i].off_width = 8;
    }
    return kc;
}

void run_build(H *h) {
    h->slot_cols_ready = false;          // the slot-ordered payload copies belong to the previous table
    h->probed_since_build = 0;
    const int64_t n = h->n_build;
    const int kw = h->table_kw;
    const float lf = h->cfg.load_factor > 0.f ? std::min(h->cfg.load_factor, 0.9f) : 0.5f;

    // ---- layout choice: direct while the table fits the L2 budget
    uint64_t direct_slots = table_slots_for(n, lf);
    // measured on B200 (profiles/r1_i_sizing_sweep.txt): the un-partitioned probe beats partition + probe up to a ~130 MB table
    const int64_t budget = h->cfg.l2_budget_bytes > 0 ? h->cfg.l2_budget_bytes : h->l2_bytes / 10 * 9;
    int32_t layout = h->cfg.layout;
    if (layout == B200J_LAYOUT_AUTO) layout = ((int64_t)direct_slots * 16 <= budget) ? B200J_LAYOUT_DIRECT : B200J_LAYOUT_RADIX;
    // A small direct table is spread thinner when the caller left the load factor open: at 0.5 about a third of the 2-slot
    // buckets are full, and a probe key that is NOT in the table then has to walk (measured: 95 % misses cost 13.3 ms per 1 B
    // probe rows against 9.1 ms for 100 % hits). Down to 0.2 while the table stays within ~48 MB, well inside the L2.
    if (layout == B200J_LAYOUT_DIRECT && h->cfg.load_factor <= 0.f && n > 0) {
        const float thin = std::max(h->thin_lf_min, std::min(0.5f, (float)((double)n * 16.0 / (double)h->thin_table_bytes)));
        if (thin < lf) direct_slots = table_slots_for(n, thin);
    }
    h->layout = layout;

    dev_alloc(h, h->zero_slot, 2 * sizeof(Slot));       // a whole bucket: the probe loads 32 B
    CU(cudaMemsetAsync(h->zero_slot.p, 0, 2 * sizeof(Slot), h->stream));
    CU(cudaMemsetAsync(h->d_ctr, 0, sizeof(Counters), h->stream));
    const int jt = h->cfg.join_type;
    if (jt == B200J_LEFT || jt == B200J_FULL || jt == B200J_LEFT_SEMI || jt == B200J_LEFT_ANTI) {
        dev_alloc(h, h->visited, (size_t)((n + 31) / 32 + 1) * 4);
        CU(cudaMemsetAsync(h->visited.p, 0, (size_t)((n + 31) / 32 + 1) * 4, h->stream));
    }
    const void *keys = h->bkeys[0].values.p;
    const uint8_t *validity = h->bkeys[0].has_nulls ? (const uint8_t *)h->bkeys[0].validity.p : nullptr;
    void *comp_hash = nullptr, *comp_ok = nullptr;      // composite keys: row hashes + joinable bitmap, needed during the build only
    struct CompGuard { H *h; void **a, **b; ~CompGuard() { if (*a) big_free(h, *a); if (*b) big_free(h, *b); } } comp_guard{h, &comp_hash, &comp_ok};
    if (h->composite && n > 0) {
        CU(big_alloc(h, &comp_hash, (size_t)n * 8 + 16));
        CU(big_alloc(h, &comp_ok, (size_t)((n + 31) / 32 + 1) * 4));
        const KeyCols kc = build_key_cols(h);
        PhaseTimer tm(h, &h->st.ms_build_insert);
        composite_hash_kernel<<<grid_for(h, n, 256, 8), 256, 0, h->stream>>>(kc, n, h->hk, (uint64_t *)comp_hash, (uint32_t *)comp_ok);
        h->st.kernel_launches++;
        CU(cudaGetLastError());
        tm.stop();
        keys = comp_hash;
        validity = (const uint8_t *)comp_ok;
    }

    BuildArgs a{};
    a.hk = h->hk;
    a.ctr = h->d_ctr;

    if (layout == B200J_LAYOUT_DIRECT || n == 0) {
        h->layout = B200J_LAYOUT_DIRECT;
        h->radix_bits = 0;
        PhaseTimer tm(h, &h->st.ms_build_insert);
        h->n_slots = direct_slots;
        dev_alloc(h, h->slots, (size_t)direct_slots * sizeof(Slot));
        CU(cudaMemsetAsync(h->slots.p, 0, (size_t)direct_slots * sizeof(Slot), h->stream));
        a.keys = keys;
        a.validity = validity;
        a.rowid = nullptr;
        a.n = n;
        a.mis = 0;      // the build store is 256 B aligned
        a.tv = h->view();
        void *slot_of = nullptr;
        struct SlotOfGuard { H *h; void **p; ~SlotOfGuard() { if (*p) big_free(h, *p); } } slot_of_guard{h, &slot_of};
        if (n > 0 && direct_slots < 0xffffffffULL) {
            CU(big_alloc(h, &slot_of, (size_t)n * 4 + 16));
            a.slot_of = (uint32_t *)slot_of;
            a.dup_bit = direct_slots < 0x7fffffffULL ? 1u : 0u;
        }
        if (n > 0) {
            const int grid = grid_for(h, n, TILE_ROWS, 8);
            launch_by_kw(kw, [&](auto KW) { build_insert_kernel<decltype(KW)::value, false><<<grid, TILE_THREADS, 0, h->stream>>>(a); });
            h->st.kernel_launches++;
        }
        CU(cudaGetLastError());
        CU(cudaMemcpyAsync(h->h_ctr, h->d_ctr, sizeof(Counters), cudaMemcpyDeviceToHost, h->stream));
        CU(cudaStreamSynchronize(h->stream));
        if (h->h_ctr->dup_rows > 0) {
            const uint64_t need = h->h_ctr->dup_rows + 2 * h->h_ctr->multi_slots + 64;
            REQUIRE(need < 0xffffffffULL, B200J_ERR_UNSUPPORTED, "duplicate-run index exceeds 32 bits");
            dev_alloc(h, h->rowids, (size_t)need * 4);
            a.tv = h->view();
            csr_scan_kernel<<<grid_for(h, (int64_t)direct_slots + 1, 256, 8), 256, 0, h->stream>>>(a.tv, direct_slots, h->d_ctr);
            const int grid = grid_for(h, n, TILE_ROWS, 8);
            launch_by_kw(kw, [&](auto KW) { csr_fill_kernel<decltype(KW)::value, false><<<grid, TILE_THREADS, 0, h->stream>>>(a); });
            h->st.kernel_launches += 2;
            CU(cudaGetLastError());
        }
        tm.stop();
    } else {
        // load factor left open: 0.42 behind the radix pass (profiles/r1_j_radix_load_factor.txt). At 100 M unique keys, 2^7
        // partitions: insert 3.66 -> 3.48 ms and probe kernel 6.97 -> 6.74 ms per 1 B rows against 0.5; 0.35 is better still for
        // probe keys that miss (11.7 -> 10.4 ms at 95 % misses) but makes duplicate-heavy builds slower (C3 exp: insert 9.3 ->
        // 11.5 ms), 0.42 is neutral there; at 0.25 the partition outgrows its share of the L2.
        radix_build(h, keys, validity, kw, n, h->cfg.load_factor > 0.f ? lf : h->radix_lf, budget);
    }
    CU(cudaStreamSynchronize(h->stream));
    h->st.layout = h->layout;
    h->st.radix_bits = h->radix_bits;
    h->st.build_rows = n;
    h->st.build_null_keys = (int64_t)h->h_ctr->null_keys;
    h->st.distinct_keys = (int64_t)h->h_ctr->distinct;
    h->st.duplicate_rows = (int64_t)h->h_ctr->dup_rows;
    h->st.table_slots = (int64_t)h->n_slots;
    h->st.table_bytes = (int64_t)h->n_slots * (int64_t)sizeof(Slot);
    h->built = true;
}

void do_build_finish(H *h, int32_t partition) {
    REQUIRE(partition >= 0 && partition < h->cfg.n_partitions, B200J_ERR_INVALID, "partition out of range");
    REQUIRE(!h->build_done[partition], B200J_ERR_STATE, "State already consumed: build_finish called twice for this partition");
    REQUIRE(!h->shuffle_armed, B200J_ERR_STATE, "build_finish between b200j_shuffle_count and b200j_shuffle_scatter");
    h->build_done[partition] = 1;
    h->n_build_done++;
    if (h->n_build_done == h->cfg.n_partitions) run_build(h);   // the last owner does the work (LimitedRc / InitializeLast)
}

// ---------------------------------------------------------------------------------------------
// probe
// ---------------------------------------------------------------------------------------------
struct ProbePlan {
    int mode;
    bool mark;
};

ProbePlan plan_for(int join_type, bool matches_only) {
    if (matches_only) return {PM_PAIRS, false};
    switch (join_type) {
        case B200J_INNER: return {PM_PAIRS, false};
        case B200J_LEFT: return {PM_PAIRS, true};
        case B200J_RIGHT: return {PM_OUTER, false};
        case B200J_FULL: return {PM_OUTER, true};
        case B200J_LEFT_SEMI:
        case B200J_LEFT_ANTI: return {PM_MARK, true};
        case B200J_RIGHT_SEMI: return {PM_SEMI, false};
        default: return {PM_ANTI, false};
    }
}

template <int KW, bool RADIX>
void launch_probe_mode(const ProbePlan &pl, int grid, cudaStream_t s, const ProbeArgs &a) {
    if (pl.mode == PM_PAIRS && !pl.mark) probe_kernel<KW, PM_PAIRS, false, RADIX><<<grid, PROBE_THREADS, 0, s>>>(a);
    else if (pl.mode == PM_PAIRS) probe_kernel<KW, PM_PAIRS, true, RADIX><<<grid, PROBE_THREADS, 0, s>>>(a);
    else if (pl.mode == PM_OUTER && !pl.mark) probe_kernel<KW, PM_OUTER, false, RADIX><<<grid, PROBE_THREADS, 0, s>>>(a);
    else if (pl.mode == PM_OUTER) probe_kernel<KW, PM_OUTER, true, RADIX><<<grid, PROBE_THREADS, 0, s>>>(a);
    else if (pl.mode == PM_SEMI) probe_kernel<KW, PM_SEMI, false, RADIX><<<grid, PROBE_THREADS, 0, s>>>(a);
    else if (pl.mode == PM_ANTI) probe_kernel<KW, PM_ANTI, false, RADIX><<<grid, PROBE_THREADS, 0, s>>>(a);
    else probe_kernel<KW, PM_MARK, true, RADIX><<<grid, PROBE_THREADS, 0, s>>>(a);
}

template <int KW, bool RADIX>
void launch_probe_fast(const ProbePlan &pl, int grid, cudaStream_t s, const ProbeArgs &a) {
    if (pl.mode == PM_PAIRS && !pl.mark) probe_fast_kernel<KW, PM_PAIRS, false, RADIX><<<grid, PF_THREADS, 0, s>>>(a);
    else if (pl.mode == PM_PAIRS) probe_fast_kernel<KW, PM_PAIRS, true, RADIX><<<grid, PF_THREADS, 0, s>>>(a);
    else if (pl.mode == PM_OUTER && !pl.mark) probe_fast_kernel<KW, PM_OUTER, false, RADIX><<<grid, PF_THREADS, 0, s>>>(a);
    else if (pl.mode == PM_OUTER) probe_fast_kernel<KW, PM_OUTER, true, RADIX><<<grid, PF_THREADS, 0, s>>>(a);
    else if (pl.mode == PM_SEMI) probe_fast_kernel<KW, PM_SEMI, false, RADIX><<<grid, PF_THREADS, 0, s>>>(a);
    else if (pl.mode == PM_ANTI) probe_fast_kernel<KW, PM_ANTI, false, RADIX><<<grid, PF_THREADS, 0, s>>>(a);
    else probe_fast_kernel<KW, PM_MARK, true, RADIX><<<grid, PF_THREADS, 0, s>>>(a);
}

void launch_probe(H *h, const ProbePlan &pl, const ProbeArgs &a, int kw, bool radix) {
    h->st.kernel_launches++;
    h->st.probe_launches++;
    if (a.mis == 0 && a.validity == nullptr && (!radix || a.rowid != nullptr)) {
        // the lean kernel: aligned keys, no validity bitmap (always the case behind the partition pass)
        const int grid = grid_for(h, a.n, PF_STAGE, h->probe_ctas_per_sm);
        if (kw == 8) {
            if (radix) launch_probe_fast<8, true>(pl, grid, h->stream, a);
            else launch_probe_fast<8, false>(pl, grid, h->stream, a);
        } else {
            if (radix) launch_probe_fast<4, true>(pl, grid, h->stream, a);
            else launch_probe_fast<4, false>(pl, grid, h->stream, a);
        }
        return;
    }
    const int grid = grid_for(h, a.n + a.mis, PROBE_STAGE, 2048 / PROBE_THREADS);
    if (kw == 8) {
        if (radix) launch_probe_mode<8, true>(pl, grid, h->stream, a);
        else launch_probe_mode<8, false>(pl, grid, h->stream, a);
    } else {
        if (radix) launch_probe_mode<4, true>(pl, grid, h->stream, a);
        else launch_probe_mode<4, false>(pl, grid, h->stream, a);
    }
}

int64_t estimate_out_cap(H *h, int64_t n_probe, const ProbePlan &pl) {
    if (pl.mode == PM_MARK) return 0;
    if (pl.mode == PM_SEMI || pl.mode == PM_ANTI) return n_probe;
    if (h->st.duplicate_rows == 0) return n_probe;      // unique build keys: at most one row per probe row
    const double avg = h->st.distinct_keys > 0 ? (double)(h->st.build_rows - h->st.build_null_keys) / (double)h->st.distinct_keys : 1.0;
    return (int64_t)((double)n_probe * avg * 1.25) + 4096 + (pl.mode == PM_OUTER ? n_probe : 0);
}

void fill_result(b200j_result *out, ResultOwner *ro, int64_t n, int mem) {
    out->mem = mem;
    out->n = n;
    out->build_idx = (uint64_t *)ro->build_idx;
    out->probe_idx = (uint32_t *)ro->probe_idx;
    out->_owner = ro;
    if (out->struct_size >= offsetof(b200j_result, build_idx_width) + sizeof(int32_t)) out->build_idx_width = ro->bwidth;
    if (out->struct_size >= offsetof(b200j_result, probe_order) + sizeof(int32_t)) out->probe_order = ro->probe_order ? 1 : 0;
}

void drop_seg(H *h, ResultOwner *ro) {
    if (ro->seg) {
        h->seg_owners.erase(ro);
        big_free(h, ro->seg);
        ro->seg = nullptr;
    }
    if (ro->aux_keys) big_free(h, ro->aux_keys);
    if (ro->aux_slot) big_free(h, ro->aux_slot);
    ro->aux_keys = ro->aux_slot = nullptr;
}

void release_owner(H *h, ResultOwner *ro) {
    if (!ro) return;
    drop_seg(h, ro);
    if (ro->mem == B200J_MEM_DEVICE) {
        if (ro->build_idx) big_free(h, ro->build_idx);
        if (ro->probe_idx) big_free(h, ro->probe_idx);
    } else {
        pinned_give(h, ro->build_idx);
        pinned_give(h, ro->probe_idx);
    }
    delete ro;
}

// Device-resident probe of `n` keys. Output arrays are device memory owned by *ro. Returns rows produced.
int64_t probe_device(H *h, const void *d_keys, const uint8_t *d_validity, int64_t bit_offset, int64_t n, const ProbePlan &pl,
                     bool count_only, ResultOwner *ro, int64_t *n_build_null) {
    const int kw = h->table_kw;
    const bool want_build = (pl.mode == PM_PAIRS || pl.mode == PM_OUTER);
    const bool radix = h->layout == B200J_LAYOUT_RADIX;
    int64_t cap = count_only ? 0 : estimate_out_cap(h, n, pl);
    *n_build_null = 0;
    if (n == 0) return 0;
    if (radix && !count_only && h->dense_ok && h->block_part_ok && h->radix_bits <= PB_MAX_BITS && h->st.duplicate_rows == 0 &&
        (pl.mode == PM_PAIRS || pl.mode == PM_OUTER))
        return probe_dense_blocked(h, d_keys, d_validity, bit_offset, n, pl, ro, n_build_null);
    std::vector<void *> temps;
    Trace tr_;
    try {
        // what the probe kernel reads: the caller's keys (direct) or the partitioned (key, row id) stream (radix)
        const void *pkeys = d_keys;
        const uint8_t *pvalid = d_validity;
        int64_t pboff = bit_offset, pn = n, n_null_rows = 0;
        const uint32_t *prid = nullptr;
        if (radix) {
            void *tk = nullptr, *tr = nullptr;
            CU(big_alloc(h, &tk, (size_t)n * kw + 16));
            temps.push_back(tk);
            CU(big_alloc(h, &tr, (size_t)n * 4 + 16));
            temps.push_back(tr);
            tr_.mark("alloc partition buffers");
            radix_partition(h, d_keys, d_validity, bit_offset, kw, n, h->radix_bits, tk, (uint32_t *)tr, &h->st.ms_probe_partition);
            tr_.mark("radix_partition");
            pkeys = tk;
            pvalid = nullptr;
            pboff = 0;
            prid = (const uint32_t *)tr;
            pn = (int64_t)h->rscratch.h_off[1 << h->radix_bits];
            n_null_rows = n - pn;
        }
        const int64_t extra = (pl.mode == PM_OUTER || pl.mode == PM_ANTI) ? n_null_rows : 0;
        // Dense path (probe.cuh): unique build keys behind the partition pass. The partitioned row-id array becomes the
        // result's probe-index column as it is; the kernel writes one build index (or NULL) per partitioned row.
        // Direct layout: same kernel over the caller's keys when they can be read as 16 B vectors and carry no bitmap.
        // Inner / Left joins pay a compaction pass when the dense result has holes: with many misses (a filtered dimension
        // table on the build side) the staged kernel, which never writes the holes, is the cheaper one on the direct layout.
        const bool dense_direct = !radix && pvalid == nullptr && ((uintptr_t)pkeys & 15) == 0 && (pl.mode == PM_OUTER || h->miss_frac < h->dense_miss_max);
        if ((radix || dense_direct) && !count_only && h->dense_ok && h->st.duplicate_rows == 0 && pn > 0 && (pl.mode == PM_PAIRS || pl.mode == PM_OUTER)) {
            void *tr = nullptr;
            if (radix) {
                tr = temps.back();                // the row-id array: owned by the result from here on
                temps.pop_back();
            } else {
                CU(big_alloc(h, &tr, (size_t)n * 4 + 16));
            }
            void *ob = nullptr;
            try {
                const int64_t rows = pl.mode == PM_OUTER ? n : pn;
                CU(big_alloc(h, &ob, (size_t)rows * 8 + 16));
                tr_.mark("alloc out_build");
                CU(cudaMemsetAsync(h->d_ctr, 0, COUNTERS_LAUNCH_BYTES, h->stream));
                ProbeArgs a{};
                a.keys = pkeys;
                a.rowid = prid;
                a.n = pn;
                a.tv = h->view();
                a.hk = h->hk;
                a.out_build = (uint64_t *)ob;
                a.out_probe = (uint32_t *)tr;
                a.out_cap = (uint64_t)rows;
                a.visited = (uint32_t *)h->visited.p;
                a.ctr = h->d_ctr;
                {
                    PhaseTimer tm(h, &h->st.ms_probe);
                    const int grid = grid_for(h, pn, PD_CHUNK * PD_WARPS, h->dense_ctas_per_sm);
                    // L1-allocating bucket loads: a small direct table, or a skewed batch behind the radix pass (a partition far
                    // above the mean size = heavy-hitter keys, whose buckets every SM would otherwise fetch from the same L2 sectors)
                    bool l1 = a.tv.l1 != 0u;
                    if (radix) {
                        const int P = 1 << h->radix_bits;
                        int64_t biggest = 0;
                        for (int p = 0; p < P; p++) biggest = std::max<int64_t>(biggest, (int64_t)h->rscratch.h_off[p + 1] - (int64_t)h->rscratch.h_off[p]);
                        l1 = h->skew_l1 && P >= 8 && biggest * P > 3 * pn;
                    }
                    auto launch = [&](auto KW, auto MARK) {
                        constexpr int kw_ = decltype(KW)::value;
                        constexpr bool mk_ = decltype(MARK)::value;
                        if (radix && l1) probe_dense_kernel<kw_, mk_, true, true><<<grid, PD_THREADS, 0, h->stream>>>(a);
                        else if (radix) probe_dense_kernel<kw_, mk_, true, false><<<grid, PD_THREADS, 0, h->stream>>>(a);
                        else if (l1) probe_dense_kernel<kw_, mk_, false, true><<<grid, PD_THREADS, 0, h->stream>>>(a);
                        else probe_dense_kernel<kw_, mk_, false, false><<<grid, PD_THREADS, 0, h->stream>>>(a);
                    };
                    if (kw == 8) {
                        if (pl.mark) launch(std::integral_constant<int, 8>{}, std::true_type{});
                        else launch(std::integral_constant<int, 8>{}, std::false_type{});
                    } else {
                        if (pl.mark) launch(std::integral_constant<int, 4>{}, std::true_type{});
                        else launch(std::integral_constant<int, 4>{}, std::false_type{});
                    }
                    h->st.kernel_launches++;
                    h->st.probe_launches++;
                    h->st.dense_probe_launches++;
                    CU(cudaGetLastError());
                    tm.stop();
                }
                CU(cudaMemcpyAsync(h->h_ctr, h->d_ctr, sizeof(Counters), cudaMemcpyDeviceToHost, h->stream));
                CU(cudaStreamSynchronize(h->stream));
                const int64_t misses = (int64_t)h->h_ctr->build_null_out;
                h->miss_frac = (float)((double)misses / (double)pn);
                tr_.mark("dense probe + counters");
                if (pl.mode == PM_OUTER) {
                    if (extra > 0) {      // NULL-key probe rows: their row ids already sit at the tail of the row-id array
                        fill_u64_kernel<<<grid_for(h, extra, 256, 8), 256, 0, h->stream>>>((uint64_t *)ob + pn, extra, ~0ULL);
                        h->st.kernel_launches++;
                        CU(cudaGetLastError());
                        CU(cudaStreamSynchronize(h->stream));
                    }
                    *n_build_null = misses + extra;
                    ro->build_idx = ob;
                    ro->probe_idx = tr;
                    free_temps(h, temps);
                    return n;
                }
                if (misses == 0) {
                    ro->build_idx = ob;
                    ro->probe_idx = tr;
                    free_temps(h, temps);
                    tr_.mark("free temps");
                    return pn;
                }
                // Inner / Left with misses: one compaction pass
                const int64_t keep = pn - misses;
                void *ob2 = nullptr, *op2 = nullptr;
                try {
                    CU(big_alloc(h, &ob2, (size_t)std::max<int64_t>(keep, 1) * 8));
                    CU(big_alloc(h, &op2, (size_t)std::max<int64_t>(keep, 1) * 4));
                    CU(cudaMemsetAsync(h->d_ctr, 0, COUNTERS_LAUNCH_BYTES, h->stream));
                    PhaseTimer tm(h, &h->st.ms_probe);
                    compact_pairs_kernel<<<grid_for(h, pn, CP_TILE, 8), CP_THREADS, 0, h->stream>>>((const uint64_t *)ob, (const uint32_t *)tr, (uint64_t)pn,
                                                                                                    (uint64_t *)ob2, (uint32_t *)op2, &h->d_ctr->out_count);
                    h->st.kernel_launches++;
                    h->st.compact_launches++;
                    CU(cudaGetLastError());
                    tm.stop();
                    CU(cudaStreamSynchronize(h->stream));
                } catch (...) {
                    if (ob2) big_free(h, ob2);
                    if (op2) big_free(h, op2);
                    throw;
                }
                big_free(h, ob);
                big_free(h, tr);
                ro->build_idx = ob2;
                ro->probe_idx = op2;
                free_temps(h, temps);
                return keep;
            } catch (...) {
                if (ob) big_free(h, ob);
                big_free(h, tr);
                throw;
            }
        }
        for (int attempt = 0; attempt < 3; attempt++) {
            void *ob = nullptr, *op = nullptr;
            if (cap > 0) {
                if (want_build) CU(big_alloc(h, &ob, (size_t)cap * 8));
                CU(big_alloc(h, &op, (size_t)cap * 4));
            }
            CU(cudaMemsetAsync(h->d_ctr, 0, COUNTERS_LAUNCH_BYTES, h->stream));
            ProbeArgs a{};
            a.keys = pkeys;
            a.validity = pvalid;
            a.bit_offset = pboff;
            a.rowid = prid;
            a.n = pn;
            a.mis = (int)(((uintptr_t)pkeys & 15) / (uintptr_t)kw);
            a.tv = h->view();
            a.hk = h->hk;
            a.out_build = (uint64_t *)ob;
            a.out_probe = (uint32_t *)op;
            a.out_cap = (uint64_t)cap;
            a.visited = (uint32_t *)h->visited.p;
            a.ctr = h->d_ctr;
            constexpr uint32_t BIG_CAP = 1u << 18;
            if (h->st.duplicate_rows >= (int64_t)BIG_RUN_ROWS && (cap > 0 || pl.mark)) {
                dev_alloc(h, h->bigruns, (size_t)BIG_CAP * sizeof(BigRun));
                a.big_runs = (BigRun *)h->bigruns.p;
                a.big_cap = BIG_CAP;
            }
            if (pn > 0) {
                PhaseTimer tm(h, &h->st.ms_probe);
                launch_probe(h, pl, a, kw, radix);
                CU(cudaGetLastError());
                tm.stop();
            }
            CU(cudaMemcpyAsync(h->h_ctr, h->d_ctr, sizeof(Counters), cudaMemcpyDeviceToHost, h->stream));
            CU(cudaStreamSynchronize(h->stream));
            const int64_t total = (int64_t)h->h_ctr->out_count;
            *n_build_null = (int64_t)h->h_ctr->build_null_out + extra;
            if (count_only || total + extra <= cap) {
                const uint32_t n_big = a.big_runs ? std::min<uint32_t>(h->h_ctr->big_count, a.big_cap) : 0u;
                if (n_big > 0) {
                    // heavy-hitter runs the probe kernel only reserved room for: the whole grid copies them (marking is
                    // idempotent, so a retried launch that deferred them twice is harmless)
                    PhaseTimer tm(h, &h->st.ms_probe);
                    const int grid = h->sm_count * 8;
                    const bool emit = want_build && !count_only;
                    if (emit && pl.mark) expand_runs_kernel<true, true><<<grid, 256, 0, h->stream>>>(a.big_runs, n_big, a.tv.rowids, a.out_build, a.out_probe, a.out_cap, a.visited);
                    else if (emit) expand_runs_kernel<true, false><<<grid, 256, 0, h->stream>>>(a.big_runs, n_big, a.tv.rowids, a.out_build, a.out_probe, a.out_cap, a.visited);
                    else if (pl.mark) expand_runs_kernel<false, true><<<grid, 256, 0, h->stream>>>(a.big_runs, n_big, a.tv.rowids, a.out_build, a.out_probe, a.out_cap, a.visited);
                    h->st.kernel_launches++;
                    CU(cudaGetLastError());
                    tm.stop();
                    CU(cudaStreamSynchronize(h->stream));
                }
                if (extra > 0 && !count_only) {
                    emit_null_rows_kernel<<<grid_for(h, extra, 256, 8), 256, 0, h->stream>>>(prid + pn, extra, (uint64_t *)ob, (uint32_t *)op,
                                                                                             (uint64_t)total);
                    h->st.kernel_launches++;
                    CU(cudaGetLastError());
                    CU(cudaStreamSynchronize(h->stream));
                }
                if (pl.mode == PM_PAIRS && h->st.duplicate_rows == 0 && pn > 0) h->miss_frac = (float)(1.0 - (double)total / (double)pn);
                ro->build_idx = ob;
                ro->probe_idx = op;
                free_temps(h, temps);
                return total + extra;
            }
            // the pair buffer was too small (duplicate keys): the kernel counted everything, run again with room.
            // Marking is idempotent, so the re-run is safe for the visited bitmap.
            if (ob) big_free(h, ob);
            if (op) big_free(h, op);
            cap = total + extra;
            h->st.probe_retries++;
        }
        throw Fail{B200J_ERR_CUDA, "probe output did not converge"};
    } catch (...) {
        free_temps(h, temps);
        throw;
    }
}

// B200J_OUT_BUILD_U32: the device result's build indices narrowed to 32 bits (in a new buffer; the wide one is recycled)
void narrow_result(H *h, ResultOwner *ro, int64_t n) {
    if (ro->bwidth == 4 || !ro->build_idx || ro->mem != B200J_MEM_DEVICE) return;
    void *nb = nullptr;
    CU(big_alloc(h, &nb, (size_t)std::max<int64_t>(n, 1) * 4 + 16));
    if (n > 0) {
        narrow_u64_u32_kernel<<<grid_for(h, n / 2 + 1, 256, 8), 256, 0, h->stream>>>((const uint64_t *)ro->build_idx, (uint32_t *)nb, n);
        h->st.kernel_launches++;
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { big_free(h, nb); CU(e); }
    big_free(h, ro->build_idx);          // stream-ordered: the kernel above has been enqueued before any reuse
    ro->build_idx = nb;
    ro->bwidth = 4;
}

// B200J_OUT_PROBE_ORDER: the device pairs of a PM_OUTER probe against unique build keys (exactly one pair per probe row, probe
// indices = base .. base + n - 1 in some order) become one build-index array in probe order; the pair arrays are recycled.
void to_probe_order(H *h, ResultOwner *ro, int64_t n, uint32_t base, bool u32) {
    if (ro->mem != B200J_MEM_DEVICE) return;
    const size_t bw = u32 ? 4 : 8;
    void *out = nullptr;
    CU(big_alloc(h, &out, (size_t)std::max<int64_t>(n, 1) * bw + 16));
    if (n > 0) {
        const int grid = grid_for(h, n, 256 * 4, 8);
        if (u32) reorder_by_probe_kernel<uint32_t><<<grid, 256, 0, h->stream>>>((const uint64_t *)ro->build_idx, (const uint32_t *)ro->probe_idx, n, base, (uint32_t *)out);
        else reorder_by_probe_kernel<unsigned long long><<<grid, 256, 0, h->stream>>>((const uint64_t *)ro->build_idx, (const uint32_t *)ro->probe_idx, n, base, (unsigned long long *)out);
        h->st.kernel_launches++;
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { big_free(h, out); CU(e); }
    drop_seg(h, ro);                          // the result no longer follows the partitioned stream
    if (ro->build_idx) big_free(h, ro->build_idx);      // stream-ordered: the kernel above has been enqueued before any reuse
    if (ro->probe_idx) big_free(h, ro->probe_idx);
    ro->build_idx = out;
    ro->probe_idx = nullptr;
    ro->bwidth = (int32_t)bw;
    ro->probe_order = true;
}

void move_result_to_host(H *h, ResultOwner *ro, int64_t n) {
    void *hb = nullptr, *hp = nullptr;
    if (ro->build_idx) {
        hb = pinned_take(h, (size_t)std::max<int64_t>(n, 1) * ro->bwidth);
        CU(cudaMemcpyAsync(hb, ro->build_idx, (size_t)n * ro->bwidth, cudaMemcpyDeviceToHost, h->stream));
    }
    if (ro->probe_idx) {
        hp = pinned_take(h, (size_t)std::max<int64_t>(n, 1) * 4);
        CU(cudaMemcpyAsync(hp, ro->probe_idx, (size_t)n * 4, cudaMemcpyDeviceToHost, h->stream));
    }
    CU(cudaStreamSynchronize(h->stream));
    if (ro->build_idx) big_free(h, ro->build_idx);
    if (ro->probe_idx) big_free(h, ro->probe_idx);
    ro->build_idx = hb;
    ro->probe_idx = hp;
    ro->mem = B200J_MEM_HOST;
    drop_seg(h, ro);
}

// ---------------------------------------------------------------------------------------------
// Host batch -> host result, pipelined (the double-buffered pinned-memory feeder).
// The batch is cut into chunks; while chunk k is partitioned and probed on the join's stream, chunk k+1 is already
// crossing PCIe on `copy_in` and the pairs of chunk k-1 are going back on `copy_out`: H2D, kernels and D2H overlap and
// the call runs at the speed of the slower PCIe direction instead of the sum of all three. Chunk results land back to
// back in ONE pinned result (row ids rebased to the caller's batch), so the caller sees the same b200j_result as from
// the unchunked path. Every chunk sweeps the whole table once (3.2 GB of HBM reads at C2: noise next to PCIe).
// ---------------------------------------------------------------------------------------------
int64_t host_chunk_rows() {
    const char *e = getenv("B200J_HOST_CHUNK_ROWS");      // read per call: tests shrink it to exercise many chunks
    const long long x = e ? atoll(e) : 0;
    return (int64_t)(x > 0 ? x : (32LL << 20));
}

struct HostPipe {               // everything the pipelined probe must give back on every exit path
    H *h;
    void *dk[2] = {nullptr, nullptr};
    void *dv[2] = {nullptr, nullptr};
    cudaEvent_t ev_in[2] = {nullptr, nullptr};
    struct Pending { ResultOwner ro; cudaEvent_t done = nullptr; };
    std::vector<Pending> pending;            // device results whose D2H copy is still in flight
    void *hb = nullptr, *hp = nullptr;       // pinned result
    explicit HostPipe(H *h_) : h(h_) {}
    void retire(size_t keep) {               // wait for the oldest copies and recycle their device buffers
        while (pending.size() > keep) {
            Pending &p = pending.front();
            if (p.done) { cudaEventSynchronize(p.done); cudaEventDestroy(p.done); }
            if (p.ro.build_idx) big_free(h, p.ro.build_idx);
            if (p.ro.probe_idx) big_free(h, p.ro.probe_idx);
            pending.erase(pending.begin());
        }
    }
    ~HostPipe() {
        h->probe_rid_base = 0;
        cudaStreamSynchronize(h->copy_in);
        cudaStreamSynchronize(h->copy_out);
        retire(0);
        for (int i = 0; i < 2; i++) {
            if (dk[i]) big_free(h, dk[i]);
            if (dv[i]) big_free(h, dv[i]);
            if (ev_in[i]) cudaEventDestroy(ev_in[i]);
        }
        if (hb) pinned_give(h, hb);
        if (hp) pinned_give(h, hp);
    }
};

int64_t probe_host_pipelined(H *h, const b200j_column &kc, int64_t n, const ProbePlan &pl, bool count_only, bool narrow, ResultOwner *ro,
                             int64_t *n_build_null, bool probe_order = false) {
    const size_t bw = narrow ? 4 : 8;        // bytes per build index in the pinned result
    const int kw = h->table_kw;
    const int64_t chunk = host_chunk_rows();
    const int64_t n_chunks = (n + chunk - 1) / chunk;
    const bool want_build = (pl.mode == PM_PAIRS || pl.mode == PM_OUTER);
    const bool want_probe = pl.mode != PM_MARK;
    HostPipe hp(h);
    for (int i = 0; i < 2; i++) {
        CU(big_alloc(h, &hp.dk[i], (size_t)chunk * kw + 16));
        if (kc.validity) CU(big_alloc(h, &hp.dv[i], (size_t)(chunk / 8) + 64));
        CU(cudaEventCreateWithFlags(&hp.ev_in[i], cudaEventDisableTiming));
    }
    // the buffers may have been used by earlier work on the join's stream: the copy stream starts after it
    cudaEvent_t ev_start = nullptr;
    CU(cudaEventCreateWithFlags(&ev_start, cudaEventDisableTiming));
    cudaEventRecord(ev_start, h->stream);
    cudaStreamWaitEvent(h->copy_in, ev_start, 0);
    cudaEventDestroy(ev_start);

    int64_t host_cap = 0;
    if (!count_only && want_probe) {
        host_cap = std::max<int64_t>(estimate_out_cap(h, n, pl), 1);
        if (want_build) hp.hb = pinned_take(h, (size_t)host_cap * bw);
        if (!probe_order) hp.hp = pinned_take(h, (size_t)host_cap * 4);      // probe order: one array, build_idx[i] of probe row i
    }
    auto issue_h2d = [&](int64_t k) {
        const int b = (int)(k & 1);
        const int64_t lo = k * chunk, m = std::min(chunk, n - lo);
        CU(cudaMemcpyAsync(hp.dk[b], (const char *)kc.values + (size_t)lo * kw, (size_t)m * kw, cudaMemcpyHostToDevice, h->copy_in));
        if (kc.validity) {
            const int64_t first = (kc.bit_offset + lo) >> 3, last = (kc.bit_offset + lo + m + 7) >> 3;
            CU(cudaMemcpyAsync(hp.dv[b], kc.validity + first, (size_t)(last - first), cudaMemcpyHostToDevice, h->copy_in));
        }
        CU(cudaEventRecord(hp.ev_in[b], h->copy_in));
    };
    int64_t total = 0, bnull = 0;
    issue_h2d(0);
    for (int64_t k = 0; k < n_chunks; k++) {
        const int b = (int)(k & 1);
        const int64_t lo = k * chunk, m = std::min(chunk, n - lo);
        // chunk k+1 goes into the other buffer, whose last reader (chunk k-1) finished before probe_device returned
        if (k + 1 < n_chunks) issue_h2d(k + 1);
        CU(cudaStreamWaitEvent(h->stream, hp.ev_in[b], 0));
        h->probe_rid_base = (uint32_t)lo;
        HostPipe::Pending pd;
        int64_t bn = 0;
        const int64_t cnt = probe_device(h, hp.dk[b], kc.validity ? (const uint8_t *)hp.dv[b] : nullptr, (kc.bit_offset + lo) & 7, m, pl,
                                         count_only, &pd.ro, &bn);
        h->probe_rid_base = 0;
        bnull += bn;
        if (count_only || !want_probe) {
            if (pd.ro.build_idx) big_free(h, pd.ro.build_idx);
            if (pd.ro.probe_idx) big_free(h, pd.ro.probe_idx);
            total += cnt;
            continue;
        }
        try {
            if (probe_order) {
                // the chunk's pairs -> its slice of the one probe-ordered array (row ids of the radix layouts already count from `lo`)
                REQUIRE(cnt == m, B200J_ERR_CUDA, "probe order: the chunk did not yield one pair per probe row");
                to_probe_order(h, &pd.ro, m, h->layout == B200J_LAYOUT_RADIX ? (uint32_t)lo : 0u, narrow);
                CU(cudaEventCreateWithFlags(&pd.done, cudaEventDisableTiming));
                CU(cudaEventRecord(pd.done, h->stream));
                CU(cudaStreamWaitEvent(h->copy_out, pd.done, 0));
                CU(cudaMemcpyAsync((char *)hp.hb + (size_t)lo * bw, pd.ro.build_idx, (size_t)m * bw, cudaMemcpyDeviceToHost, h->copy_out));
                CU(cudaEventRecord(pd.done, h->copy_out));
            } else {
            if (cnt > 0 && lo > 0 && h->layout != B200J_LAYOUT_RADIX) {
                add_base_u32_kernel<<<grid_for(h, cnt, 256, 8), 256, 0, h->stream>>>((uint32_t *)pd.ro.probe_idx, cnt, (uint32_t)lo);
                h->st.kernel_launches++;
                CU(cudaGetLastError());
            }
            if (total + cnt > host_cap) {
                // duplicate build keys made the result larger than estimated: grow the pinned result (rare)
                CU(cudaStreamSynchronize(h->copy_out));
                const int64_t ncap = std::max<int64_t>((int64_t)((double)(total + cnt) * ((double)n / (double)(lo + m)) * 1.1) + 4096, total + cnt);
                if (want_build) {
                    void *nb = pinned_take(h, (size_t)ncap * bw);
                    memcpy(nb, hp.hb, (size_t)total * bw);
                    pinned_give(h, hp.hb);
                    hp.hb = nb;
                }
                void *np = pinned_take(h, (size_t)ncap * 4);
                memcpy(np, hp.hp, (size_t)total * 4);
                pinned_give(h, hp.hp);
                hp.hp = np;
                host_cap = ncap;
            }
            if (narrow && want_build && cnt > 0) narrow_result(h, &pd.ro, cnt);
            // results of probe_device are complete (it synchronises the stream); a kernel may still be rebasing row ids / narrowing
            CU(cudaEventCreateWithFlags(&pd.done, cudaEventDisableTiming));
            CU(cudaEventRecord(pd.done, h->stream));
            CU(cudaStreamWaitEvent(h->copy_out, pd.done, 0));
            if (cnt > 0) {
                if (want_build)
                    CU(cudaMemcpyAsync((char *)hp.hb + (size_t)total * bw, pd.ro.build_idx, (size_t)cnt * bw, cudaMemcpyDeviceToHost, h->copy_out));
                CU(cudaMemcpyAsync((char *)hp.hp + (size_t)total * 4, pd.ro.probe_idx, (size_t)cnt * 4, cudaMemcpyDeviceToHost, h->copy_out));
            }
            CU(cudaEventRecord(pd.done, h->copy_out));
            }
        } catch (...) {
            if (pd.done) cudaEventDestroy(pd.done);
            if (pd.ro.build_idx) big_free(h, pd.ro.build_idx);
            if (pd.ro.probe_idx) big_free(h, pd.ro.probe_idx);
            throw;
        }
        hp.pending.push_back(pd);
        total += cnt;
        hp.retire(2);                 // at most two chunk results wait for PCIe: the loop runs at the speed of the D2H copies
    }
    CU(cudaStreamSynchronize(h->copy_out));
    hp.retire(0);
    *n_build_null = bnull;
    if (!count_only && want_probe) {
        ro->build_idx = hp.hb;
        ro->probe_idx = hp.hp;
        ro->mem = B200J_MEM_HOST;
        ro->bwidth = (int32_t)bw;
        ro->probe_order = probe_order;
        hp.hb = hp.hp = nullptr;
    }
    return total;
}

void do_apply_pairs(H *h, int32_t partition, int64_t n_probe_rows, int32_t mem, int64_t n_pairs, const uint64_t *build_idx,
                    const uint32_t *probe_idx, uint32_t want, b200j_result *out);

// Composite keys (n_keys > 1), the reference's own order of operations on the device:
//   calculate_hash over all key columns (shared.rs:11-16)            composite_hash_kernel
//   get_matching_indices on the HASH (shared.rs:29-47)               probe_device on the row hashes, matches only
//   equal_rows_arr on the candidate pairs (datafusion_private.rs)    equal_rows_kernel + compact_pairs_kernel
//   join-type index algebra (datafusion_private.rs:85-185)           the same code path b200j_apply_pairs uses
void do_probe_composite(H *h, int32_t partition, const b200j_batch *b, uint32_t want, b200j_result *out) {
    const int64_t n = b->n_rows;
    const int nk = h->cfg.n_keys;
    std::vector<void *> temps;
    ResultOwner cand;
    try {
        const KeyCols pc = stage_key_cols(h, b, n, temps);
        int64_t cnt = 0, bn = 0;
        if (n > 0) {
            void *dh = nullptr, *dok = nullptr;
            CU(big_alloc(h, &dh, (size_t)n * 8 + 16));
            temps.push_back(dh);
            CU(big_alloc(h, &dok, (size_t)((n + 31) / 32 + 1) * 4));
            temps.push_back(dok);
            {
                PhaseTimer tm(h, &h->st.ms_probe);
                composite_hash_kernel<<<grid_for(h, n, 256, 8), 256, 0, h->stream>>>(pc, n, h->hk, (uint64_t *)dh, (uint32_t *)dok);
                h->st.kernel_launches++;
                CU(cudaGetLastError());
                tm.stop();
            }
            cnt = probe_device(h, dh, (const uint8_t *)dok, 0, n, ProbePlan{PM_PAIRS, false}, false, &cand, &bn);
            if (cnt > 0) {
                // equality on the key VALUES: drops hash collisions
                CU(cudaMemsetAsync(h->d_ctr, 0, COUNTERS_LAUNCH_BYTES, h->stream));
                const KeyCols bc = build_key_cols(h);
                equal_rows_kernel<<<grid_for(h, cnt, 256, 8), 256, 0, h->stream>>>(bc, pc, (uint64_t *)cand.build_idx, (const uint32_t *)cand.probe_idx, cnt,
                                                                                   &h->d_ctr->build_null_out);
                h->st.kernel_launches++;
                CU(cudaGetLastError());
                CU(cudaMemcpyAsync(h->h_ctr, h->d_ctr, sizeof(Counters), cudaMemcpyDeviceToHost, h->stream));
                CU(cudaStreamSynchronize(h->stream));
                const int64_t failed = (int64_t)h->h_ctr->build_null_out;
                if (failed > 0) {
                    const int64_t keep = cnt - failed;
                    void *ob2 = nullptr, *op2 = nullptr;
                    CU(big_alloc(h, &ob2, (size_t)std::max<int64_t>(keep, 1) * 8));
                    temps.push_back(ob2);
                    CU(big_alloc(h, &op2, (size_t)std::max<int64_t>(keep, 1) * 4));
                    temps.push_back(op2);
                    CU(cudaMemsetAsync(h->d_ctr, 0, COUNTERS_LAUNCH_BYTES, h->stream));
                    compact_pairs_kernel<<<grid_for(h, cnt, CP_TILE, 8), CP_THREADS, 0, h->stream>>>((const uint64_t *)cand.build_idx, (const uint32_t *)cand.probe_idx,
                                                                                                     (uint64_t)cnt, (uint64_t *)ob2, (uint32_t *)op2, &h->d_ctr->out_count);
                    h->st.kernel_launches++;
                    h->st.compact_launches++;
                    CU(cudaGetLastError());
                    std::swap(cand.build_idx, ob2);      // the unfiltered arrays are now the temps
                    std::swap(cand.probe_idx, op2);
                    temps[temps.size() - 2] = ob2;
                    temps[temps.size() - 1] = op2;
                    cnt = keep;
                }
            }
        }
        h->st.probe_rows += n;
        if (want & B200J_OUT_MATCHES_ONLY) {
            // the caller applies a JoinFilter and comes back through b200j_apply_pairs
            ResultOwner *ro = new ResultOwner(cand);
            cand = ResultOwner{};
            try {
                if (want & B200J_OUT_COUNT_ONLY) {
                    if (ro->build_idx) big_free(h, ro->build_idx);
                    if (ro->probe_idx) big_free(h, ro->probe_idx);
                    ro->build_idx = ro->probe_idx = nullptr;
                } else {
                    if (want & B200J_OUT_BUILD_U32) narrow_result(h, ro, cnt);
                    if (want & B200J_OUT_HOST) move_result_to_host(h, ro, cnt);
                    else if (want & B200J_OUT_BUILD_U32) CU(cudaStreamSynchronize(h->stream));
                }
            } catch (...) {
                release_owner(h, ro);
                throw;
            }
            fill_result(out, ro, cnt, ro->mem);
            out->n_build_null = 0;
            out->n_probe_null = 0;
            h->st.result_rows += cnt;
        } else {
            do_apply_pairs(h, partition, n, B200J_MEM_DEVICE, cnt, (const uint64_t *)cand.build_idx, (const uint32_t *)cand.probe_idx, want, out);
        }
        if (cand.build_idx) big_free(h, cand.build_idx);
        if (cand.probe_idx) big_free(h, cand.probe_idx);
        cand = ResultOwner{};
        free_temps(h, temps);
    } catch (...) {
        if (cand.build_idx) big_free(h, cand.build_idx);
        if (cand.probe_idx) big_free(h, cand.probe_idx);
        free_temps(h, temps);
        throw;
    }
}

void do_probe_batch(H *h, int32_t partition, const b200j_batch *b, uint32_t want, b200j_result *out) {
    REQUIRE(partition >= 0 && partition < h->cfg.n_partitions, B200J_ERR_INVALID, "partition out of range");
    REQUIRE(h->built, B200J_ERR_INVALID, "probe before every partition finished the build side");
    REQUIRE(!h->probe_done[partition], B200J_ERR_STATE, "probe_batch after probe_finish");
    // the pending scatter's offsets live in the handle's partition scratch, which a radix probe would overwrite
    REQUIRE(!h->shuffle_armed, B200J_ERR_STATE, "probe_batch between b200j_shuffle_count and b200j_shuffle_scatter");
    REQUIRE(out != nullptr && out->struct_size >= RESULT_V1_SIZE, B200J_ERR_INVALID, "result struct missing or too small");
    check_batch(h, b, h->cfg.n_probe_payload);
    REQUIRE(b->n_rows < 0xffff0000LL, B200J_ERR_UNSUPPORTED, "probe batch of more than 2^32 - 2^16 rows (probe indices are UInt32)");
    if (h->composite) {
        do_probe_composite(h, partition, b, want, out);
        return;
    }
    ProbePlan pl = plan_for(h->cfg.join_type, (want & B200J_OUT_MATCHES_ONLY) != 0);
    const bool count_only = (want & B200J_OUT_COUNT_ONLY) != 0;
    const int kw = h->key_widths[0];
    const int64_t n = b->n_rows;
    // B200J_OUT_PROBE_ORDER: unique build keys -> every probe row has at most one match; probed as an outer join (one pair per
    // probe row, NULL build index where there is none) and handed back as one array in the order of the batch
    const int jt_ = h->cfg.join_type;
    const bool probe_order = (want & B200J_OUT_PROBE_ORDER) != 0 && !(want & (B200J_OUT_MATCHES_ONLY | B200J_OUT_COUNT_ONLY)) &&
                             out->struct_size >= offsetof(b200j_result, probe_order) + sizeof(int32_t) && h->st.duplicate_rows == 0 &&
                             kw != B200J_T_UTF8 && (jt_ == B200J_INNER || jt_ == B200J_LEFT || jt_ == B200J_RIGHT || jt_ == B200J_FULL);
    if (probe_order) pl.mode = PM_OUTER;
    ResultOwner *ro = new ResultOwner();
    ro->want_seg = !probe_order;
    ro->want_aux = ((want & B200J_OUT_GATHER_HINT) != 0 || (h->auto_hint && h->gather_seen)) && !(want & B200J_OUT_HOST) && !count_only && h->gather_seg && !probe_order;
    std::vector<void *> temps;
    try {
        int64_t total = 0, n_bnull = 0;
        if (b->mem == B200J_MEM_DEVICE || n == 0) {
            const b200j_column &kc = b->keys[0];
            total = probe_device(h, n ? kc.values : nullptr, n ? kc.validity : nullptr, n ? kc.bit_offset : 0, n, pl, count_only, ro, &n_bnull);
        } else if (((want & B200J_OUT_HOST) || count_only) && n > host_chunk_rows() + host_chunk_rows() / 2) {
            total = probe_host_pipelined(h, b->keys[0], n, pl, count_only, (want & B200J_OUT_BUILD_U32) != 0, ro, &n_bnull, probe_order);
        } else {
            const b200j_column &kc = b->keys[0];
            void *dk = nullptr;
            CU(big_alloc(h, &dk, (size_t)n * kw + 16));
            temps.push_back(dk);
            CU(cudaMemcpyAsync(dk, kc.values, (size_t)n * kw, cudaMemcpyHostToDevice, h->stream));
            int64_t boff = 0;
            const uint8_t *dv = stage_validity(h, kc, n, B200J_MEM_HOST, temps, &boff);
            total = probe_device(h, dk, dv, boff, n, pl, count_only, ro, &n_bnull);
        }
        if (count_only) {
            if (ro->build_idx) big_free(h, ro->build_idx);
            if (ro->probe_idx) big_free(h, ro->probe_idx);
            ro->build_idx = ro->probe_idx = nullptr;
        } else {
            if (probe_order && ro->mem == B200J_MEM_DEVICE) {
                REQUIRE(total == n, B200J_ERR_CUDA, "probe order: the batch did not yield one pair per probe row");
                to_probe_order(h, ro, n, 0u, (want & B200J_OUT_BUILD_U32) != 0);
                if (!(want & B200J_OUT_HOST)) CU(cudaStreamSynchronize(h->stream));
            }
            if (want & B200J_OUT_BUILD_U32) narrow_result(h, ro, total);
            if ((want & B200J_OUT_HOST) && ro->mem != B200J_MEM_HOST) move_result_to_host(h, ro, total);
            else if (want & B200J_OUT_BUILD_U32) CU(cudaStreamSynchronize(h->stream));
        }
        if (probe_order) h->st.probe_order_results++;
        free_temps(h, temps);
        fill_result(out, ro, total, ro->mem);
        out->n_build_null = n_bnull;
        out->n_probe_null = 0;
        h->st.probe_rows += n;
        h->st.result_rows += total;
    } catch (...) {
        free_temps(h, temps);
        release_owner(h, ro);
        throw;
    }
}

void do_probe_push(H *h, int32_t partition, const b200j_batch *b);
void do_probe_flush(H *h, int32_t partition, uint32_t want, b200j_result *out);
void do_gather_staged(H *h, int32_t partition, const b200j_result *pairs, uint32_t want, b200j_gathered *out);
void do_probe_pieces(H *h, int32_t partition, int32_t n_pieces, const void *const *piece_keys, const int64_t *piece_rows, int32_t piece_key_width,
                     uint32_t want, b200j_result *out);

// bitmap -> indices on the device; returns count. out64/out32: which side receives the indices.
int64_t select_bits(H *h, const uint32_t *bm, int64_t n, bool want_set, void **out64, void **out32) {
    if (out64) *out64 = nullptr;
    if (out32) *out32 = nullptr;
    if (n == 0) return 0;
    void *o64 = nullptr, *o32 = nullptr;
    try {
        if (out64) CU(big_alloc(h, &o64, (size_t)n * 8));
        if (out32) CU(big_alloc(h, &o32, (size_t)n * 4));
        CU(cudaMemsetAsync(h->d_ctr, 0, COUNTERS_LAUNCH_BYTES, h->stream));
        bitmap_select_kernel<<<grid_for(h, (n + 31) / 32, 256, 8), 256, 0, h->stream>>>(bm, n, want_set ? 1 : 0, (uint64_t *)o64, (uint32_t *)o32,
                                                                                         (uint64_t)n, &h->d_ctr->out_count);
        h->st.kernel_launches++;
        CU(cudaGetLastError());
        CU(cudaMemcpyAsync(h->h_ctr, h->d_ctr, sizeof(Counters), cudaMemcpyDeviceToHost, h->stream));
        CU(cudaStreamSynchronize(h->stream));
    } catch (...) {
        if (o64) big_free(h, o64);
        if (o32) big_free(h, o32);
        throw;
    }
    if (out64) *out64 = o64;
    if (out32) *out32 = o32;
    return (int64_t)h->h_ctr->out_count;
}

void fill_null_side(H *h, ResultOwner *ro, int64_t n, bool null_build, bool null_probe) {
    if (n == 0) return;
    if (null_build && !ro->build_idx) {
        CU(big_alloc(h, &ro->build_idx, (size_t)n * 8));
        fill_u64_kernel<<<grid_for(h, n, 256, 8), 256, 0, h->stream>>>((uint64_t *)ro->build_idx, n, ~0ULL);
        h->st.kernel_launches++;
    }
    if (null_probe && !ro->probe_idx) {
        CU(big_alloc(h, &ro->probe_idx, (size_t)n * 4));
        fill_u32_kernel<<<grid_for(h, n, 256, 8), 256, 0, h->stream>>>((uint32_t *)ro->probe_idx, n, 0xffffffffu);
        h->st.kernel_launches++;
    }
}

void do_probe_finish(H *h, int32_t partition, uint32_t want, b200j_result *out) {
    REQUIRE(partition >= 0 && partition < h->cfg.n_partitions, B200J_ERR_INVALID, "partition out of range");
    REQUIRE(h->built, B200J_ERR_INVALID, "probe_finish before the build side finished");
    REQUIRE(!h->probe_done[partition], B200J_ERR_STATE, "State already consumed: probe_finish called twice for this partition");
    REQUIRE(out != nullptr && out->struct_size >= RESULT_V1_SIZE, B200J_ERR_INVALID, "result struct missing or too small");
    h->probe_done[partition] = 1;
    h->n_probe_done++;
    ResultOwner *ro = new ResultOwner();
    try {
        int64_t n = 0;
        const int jt = h->cfg.join_type;
        const bool last = h->n_probe_done == h->cfg.n_partitions;
        const bool left_kind = jt == B200J_LEFT || jt == B200J_FULL || jt == B200J_LEFT_SEMI || jt == B200J_LEFT_ANTI;
        if (last && left_kind && h->n_build > 0) {
            // emit_unmatched_build_records (left_outer.rs:174-195, full.rs:181-201, left_anti.rs:166-179),
            // emit_matched_build_records (left_semi.rs:166-179)
            n = select_bits(h, (const uint32_t *)h->visited.p, h->n_build, jt == B200J_LEFT_SEMI, &ro->build_idx, nullptr);
            if (!(want & B200J_OUT_COUNT_ONLY)) fill_null_side(h, ro, n, false, true);
        }
        if (want & B200J_OUT_COUNT_ONLY) {
            if (ro->build_idx) big_free(h, ro->build_idx);
            ro->build_idx = nullptr;
        } else {
            if (want & B200J_OUT_BUILD_U32) narrow_result(h, ro, n);
            if (want & B200J_OUT_HOST) move_result_to_host(h, ro, n);
            else CU(cudaStreamSynchronize(h->stream));
        }
        fill_result(out, ro, n, ro->mem);
        out->n_build_null = 0;
        out->n_probe_null = n;
        h->st.result_rows += n;
    } catch (...) {
        release_owner(h, ro);
        throw;
    }
}

void do_apply_pairs(H *h, int32_t partition, int64_t n_probe_rows, int32_t mem, int64_t n_pairs, const uint64_t *build_idx,
                    const uint32_t *probe_idx, uint32_t want, b200j_result *out) {
    REQUIRE(partition >= 0 && partition < h->cfg.n_partitions, B200J_ERR_INVALID, "partition out of range");
    REQUIRE(h->built, B200J_ERR_INVALID, "apply_pairs before the build side finished");
    REQUIRE(out != nullptr && out->struct_size >= RESULT_V1_SIZE, B200J_ERR_INVALID, "result struct missing or too small");
    REQUIRE(n_pairs >= 0 && n_probe_rows >= 0 && n_probe_rows < 0xffffffffLL, B200J_ERR_INVALID, "bad pair / row counts");
    REQUIRE(n_pairs == 0 || (build_idx && probe_idx), B200J_ERR_INVALID, "pair arrays missing");
    const int jt = h->cfg.join_type;
    const bool mark = jt == B200J_LEFT || jt == B200J_FULL || jt == B200J_LEFT_SEMI || jt == B200J_LEFT_ANTI;
    const bool need_seen = jt == B200J_RIGHT || jt == B200J_FULL || jt == B200J_RIGHT_SEMI || jt == B200J_RIGHT_ANTI;
    const bool keep_pairs = jt == B200J_INNER || jt == B200J_LEFT || jt == B200J_RIGHT || jt == B200J_FULL;
    ResultOwner *ro = new ResultOwner();
    std::vector<void *> temps;
    try {
        const uint64_t *db = build_idx;
        const uint32_t *dp = probe_idx;
        if (mem == B200J_MEM_HOST && n_pairs > 0) {
            void *a = nullptr, *b = nullptr;
            CU(big_alloc(h, &a, (size_t)n_pairs * 8));
            temps.push_back(a);
            CU(big_alloc(h, &b, (size_t)n_pairs * 4));
            temps.push_back(b);
            CU(cudaMemcpyAsync(a, build_idx, (size_t)n_pairs * 8, cudaMemcpyHostToDevice, h->stream));
            CU(cudaMemcpyAsync(b, probe_idx, (size_t)n_pairs * 4, cudaMemcpyHostToDevice, h->stream));
            db = (const uint64_t *)a;
            dp = (const uint32_t *)b;
        }
        uint32_t *seen = nullptr;
        if (need_seen && n_probe_rows > 0) {
            void *s = nullptr;
            const size_t sb = (size_t)((n_probe_rows + 31) / 32 + 1) * 4;
            CU(big_alloc(h, &s, sb));
            temps.push_back(s);
            CU(cudaMemsetAsync(s, 0, sb, h->stream));
            seen = (uint32_t *)s;
        }
        if (n_pairs > 0) {
            // the pairs come from the caller (the JoinFilter round trip): an index outside the probe batch / the build side is
            // B200J_ERR_INVALID, never a write outside the bitmaps
            CU(cudaMemsetAsync(h->d_ctr, 0, COUNTERS_LAUNCH_BYTES, h->stream));
            mark_pairs_kernel<<<grid_for(h, n_pairs, 256, 8), 256, 0, h->stream>>>(db, dp, n_pairs, seen, mark ? (uint32_t *)h->visited.p : nullptr,
                                                                                   (uint64_t)n_probe_rows, (uint64_t)h->n_build, &h->d_ctr->build_null_out);
            h->st.kernel_launches++;
            CU(cudaGetLastError());
            CU(cudaMemcpyAsync(h->h_ctr, h->d_ctr, sizeof(Counters), cudaMemcpyDeviceToHost, h->stream));
            CU(cudaStreamSynchronize(h->stream));
            REQUIRE(h->h_ctr->build_null_out == 0, B200J_ERR_INVALID, "apply_pairs: pair index outside the probe batch / the build side");
        }
        int64_t n_extra = 0;
        void *extra32 = nullptr;
        if (need_seen && n_probe_rows > 0) {
            const bool want_set = jt == B200J_RIGHT_SEMI;
            n_extra = select_bits(h, seen, n_probe_rows, want_set, nullptr, &extra32);
            temps.push_back(extra32);
        }
        const int64_t n_keep = keep_pairs ? n_pairs : 0;
        const int64_t n = n_keep + n_extra;
        if (n > 0 && !(want & B200J_OUT_COUNT_ONLY)) {
            CU(big_alloc(h, &ro->build_idx, (size_t)n * 8));
            CU(big_alloc(h, &ro->probe_idx, (size_t)n * 4));
            if (n_keep) {
                CU(cudaMemcpyAsync(ro->build_idx, db, (size_t)n_keep * 8, cudaMemcpyDeviceToDevice, h->stream));
                CU(cudaMemcpyAsync(ro->probe_idx, dp, (size_t)n_keep * 4, cudaMemcpyDeviceToDevice, h->stream));
            }
            if (n_extra) {
                fill_u64_kernel<<<grid_for(h, n_extra, 256, 8), 256, 0, h->stream>>>((uint64_t *)ro->build_idx + n_keep, n_extra, ~0ULL);
                h->st.kernel_launches++;
                CU(cudaMemcpyAsync((uint32_t *)ro->probe_idx + n_keep, extra32, (size_t)n_extra * 4, cudaMemcpyDeviceToDevice, h->stream));
            }
        }
        CU(cudaGetLastError());
        if (!(want & B200J_OUT_COUNT_ONLY) && (want & B200J_OUT_BUILD_U32)) narrow_result(h, ro, n);
        if (!(want & B200J_OUT_COUNT_ONLY) && (want & B200J_OUT_HOST)) move_result_to_host(h, ro, n);
        else CU(cudaStreamSynchronize(h->stream));
        free_temps(h, temps);
        fill_result(out, ro, n, ro->mem);
        out->n_build_null = n_extra;
        out->n_probe_null = 0;
        h->st.result_rows += n;
    } catch (...) {
        free_temps(h, temps);
        release_owner(h, ro);
        throw;
    }
}

void do_hash_keys(H *h, const b200j_batch *b, uint64_t *out) {
    check_batch(h, b, 0);
    REQUIRE(out != nullptr || b->n_rows == 0, B200J_ERR_INVALID, "hash output buffer is NULL");
    REQUIRE(h->cfg.n_keys <= MAX_KEYS, B200J_ERR_UNSUPPORTED, "too many key columns");
    const int64_t n = b->n_rows;
    if (n == 0) return;
    std::vector<void *> temps;
    try {
        const KeyCols kc = stage_key_cols(h, b, n, temps);
        uint64_t *d_out = out;
        if (b->mem == B200J_MEM_HOST) {
            void *d = nullptr;
            CU(big_alloc(h, &d, (size_t)n * 8));
            temps.push_back(d);
            d_out = (uint64_t *)d;
        }
        hash_keys_kernel<<<grid_for(h, n, 256, 8), 256, 0, h->stream>>>(kc, n, h->hk, d_out);
        h->st.kernel_launches++;
        CU(cudaGetLastError());
        if (b->mem == B200J_MEM_HOST) CU(cudaMemcpyAsync(out, d_out, (size_t)n * 8, cudaMemcpyDeviceToHost, h->stream));
        CU(cudaStreamSynchronize(h->stream));
        free_temps(h, temps);
    } catch (...) {
        free_temps(h, temps);
        throw;
    }
}

void do_shuffle_count(H *h, const b200j_batch *b, int32_t n_dest, int64_t *counts);
void do_shuffle_set_hot(H *h, const uint64_t *keys, int32_t n);
void do_shuffle_scatter(H *h, int32_t n_dest, void *const *dest_keys, uint32_t *const *dest_rowid, const int64_t *dest_offset, uint32_t rowid_base);
void do_shuffle_local_scatter(H *h, const b200j_batch *b, int32_t n_dest, void *send_keys, uint32_t *send_rowid, int64_t region_rows,
                              uint32_t rowid_base, int32_t key_out_width, int64_t *counts);

template <typename F>
int32_t guarded(b200j_handle *h, F &&f) {
    if (!h) {
        g_last_error = "handle is NULL";
        return B200J_ERR_INVALID;
    }
    std::lock_guard<std::mutex> g(h->mu);
    try {
        int cur = -1;
        if (cudaGetDevice(&cur) != cudaSuccess || cur != h->cfg.device) {
            if (cudaSetDevice(h->cfg.device) != cudaSuccess) {
                cudaGetLastError();
                h->err = "cudaSetDevice failed (no CUDA device?)";
                return B200J_ERR_CUDA;
            }
        }
        f();
        return B200J_OK;
    } catch (const Fail &e) {
        h->err = e.msg;
        return e.code;
    } catch (const std::bad_alloc &) {
        h->err = "host allocation failed";
        return B200J_ERR_OOM;
    } catch (...) {
        h->err = "unknown internal error";
        return B200J_ERR_CUDA;
    }
}

}  // namespace

namespace {
#include "radix_host.inl"
#include "gather_host.inl"
#include "gather_remote_host.inl"
}  // namespace

// ---- allocation-free destination count (b200j_shuffle_count_static; no counterpart in the reference) ------------------------
// part_hist_kernel<KW, DEST> with its per-CTA histograms (the scratch h->rscratch.cta_hist of the allocating path) in a
// file-scope device array instead. A CTA takes ST_TILES_PER_CTA scatter tiles (32768 rows) whatever the batch size, so the
// array's ST_MAX_GRID columns bound the batch: B200J_STATIC_MAX_ROWS rows (+ one column for a batch that starts off a
// 16-byte boundary). Layout as everywhere: [bin][grid], bins = n_dest ranks + the NULL-key bin.
namespace {
constexpr int ST_TILES_PER_CTA = 4;
constexpr int ST_MAX_GRID = 4096;
static_assert((int64_t)ST_MAX_GRID * ST_TILES_PER_CTA * SC_TILE == (int64_t)B200J_STATIC_MAX_ROWS, "the header states the array's capacity");
static_assert(B200J_MAX_DEST <= RP_MAX_DEST, "bins of the static histogram");
constexpr size_t ST_HIST_ENTRIES = (size_t)(B200J_MAX_DEST + 1) * (ST_MAX_GRID + 1);

static __device__ __align__(256) uint32_t g_static_hist[ST_HIST_ENTRIES];       // 272 KiB on every device that loads the library

// In host code `g_static_hist` is only the host-side stand-in of the symbol, not a device address: the address the kernels
// get is looked up here, once per handle, with the handle's device current.
uint32_t *static_hist_address() {
    void *p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_static_hist);
    if (e != cudaSuccess || p == nullptr) {
        cudaGetLastError();
        throw Fail{B200J_ERR_CUDA, std::string("cudaGetSymbolAddress(g_static_hist) failed: ") + cudaGetErrorString(e)};
    }
    return (uint32_t *)p;
}

// One CTA per bin: sum of the bin's row of per-CTA counts -> counts[bin] (the caller's device buffer).
__global__ void __launch_bounds__(256) static_count_sum_kernel(const uint32_t *__restrict__ cta_hist, uint32_t grid, long long *__restrict__ counts) {
    __shared__ unsigned long long s_w[8];
    const uint32_t *rowp = cta_hist + (size_t)blockIdx.x * grid;
    unsigned long long sum = 0;
    for (uint32_t i = threadIdx.x; i < grid; i += 256) sum += rowp[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_down_sync(0xffffffffu, sum, o);
    if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = sum;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long tot = 0;
        for (int w = 0; w < 8; w++) tot += s_w[w];
        counts[blockIdx.x] = (long long)tot;
    }
}

void do_shuffle_count_static(H *h, const void *keys, const uint8_t *validity, int64_t bit_offset, int64_t n_rows, int32_t n_dest, int64_t *counts) {
    REQUIRE(h->cfg.n_keys == 1 && h->key_widths[0] != B200J_T_UTF8, B200J_ERR_UNSUPPORTED, "the exchange partitions one fixed-width key column");
    REQUIRE(n_dest >= 1 && n_dest <= B200J_MAX_DEST, B200J_ERR_INVALID, "n_dest must be in 1..16");
    REQUIRE(counts != nullptr, B200J_ERR_INVALID, "counts is NULL");
    REQUIRE(n_rows >= 0, B200J_ERR_INVALID, "n_rows is negative");
    REQUIRE(n_rows <= (int64_t)B200J_STATIC_MAX_ROWS, B200J_ERR_INVALID,
            "n_rows exceeds B200J_STATIC_MAX_ROWS (134217728): the static scratch array has no room for more; split the batch and add the counts");
    REQUIRE(n_rows == 0 || keys != nullptr, B200J_ERR_INVALID, "keys is NULL");
    REQUIRE(h->hot_n == 0, B200J_ERR_STATE, "b200j_shuffle_count_static does not route hot keys: clear the hot set or use b200j_shuffle_count");
    REQUIRE(h->static_hist != nullptr, B200J_ERR_STATE, "static scratch address missing");
    const int kw = h->key_widths[0];
    REQUIRE(((uintptr_t)keys % (uintptr_t)kw) == 0, B200J_ERR_INVALID, "keys must be aligned to the key width");
    if (n_rows == 0) {
        CU(cudaMemsetAsync(counts, 0, (size_t)(n_dest + 1) * 8, h->stream));
        return;
    }
    PartArgs a{};
    a.keys = keys;
    a.validity = validity;
    a.bit_offset = bit_offset;
    a.n = (uint32_t)n_rows;
    a.mis = (uint32_t)(((uintptr_t)keys & 15) / (uintptr_t)kw);
    a.n_dest = n_dest;
    a.n_bins = (uint32_t)n_dest;
    a.hk = h->hk;
    a.tiles_per_cta = ST_TILES_PER_CTA;
    a.cta_hist = h->static_hist;
    const int64_t n_tiles = (n_rows + a.mis + SC_TILE - 1) / SC_TILE;
    const int grid = (int)((n_tiles + ST_TILES_PER_CTA - 1) / ST_TILES_PER_CTA);
    REQUIRE(grid >= 1 && grid <= ST_MAX_GRID + 1, B200J_ERR_INVALID, "n_rows exceeds B200J_STATIC_MAX_ROWS");
    const int P1 = n_dest + 1;
    const size_t hist_smem = (size_t)P1 * 4;
    if (kw == 8) part_hist_kernel<8, true><<<grid, RP_THREADS, hist_smem, h->stream>>>(a);
    else part_hist_kernel<4, true><<<grid, RP_THREADS, hist_smem, h->stream>>>(a);
    static_count_sum_kernel<<<P1, 256, 0, h->stream>>>(h->static_hist, (uint32_t)grid, (long long *)counts);
    h->st.kernel_launches += 2;
    CU(cudaGetLastError());
}
}  // namespace

extern "C" {

int32_t b200j_abi_version(void) { return B200J_ABI_VERSION; }

int32_t b200j_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

int32_t b200j_create(const b200j_config *cfg, b200j_handle **out) {
    if (!cfg || !out) {
        g_last_error = "b200j_create: NULL argument";
        return B200J_ERR_INVALID;
    }
    *out = nullptr;
    b200j_handle *h = nullptr;
    try {
        REQUIRE(cfg->struct_size >= sizeof(b200j_config), B200J_ERR_INVALID, "config.struct_size too small (ABI mismatch)");
        REQUIRE(cfg->join_type >= 0 && cfg->join_type <= 7, B200J_ERR_INVALID, "unknown join type");
        REQUIRE(cfg->n_partitions >= 1, B200J_ERR_INVALID, "n_partitions must be >= 1");
        REQUIRE(cfg->n_keys >= 1 && cfg->n_keys <= MAX_KEYS, B200J_ERR_INVALID, "n_keys must be in 1..8");
        REQUIRE(cfg->key_widths != nullptr, B200J_ERR_INVALID, "key_widths is NULL");
        for (int i = 0; i < cfg->n_keys; i++)
            REQUIRE(cfg->key_widths[i] == 4 || cfg->key_widths[i] == 8 || cfg->key_widths[i] == B200J_T_UTF8, B200J_ERR_UNSUPPORTED,
                    "key columns must be 4 or 8 bytes wide, or B200J_T_UTF8");
        REQUIRE(cfg->n_build_payload >= 0 && cfg->n_probe_payload >= 0, B200J_ERR_INVALID, "negative payload count");
        REQUIRE(cfg->n_build_payload == 0 || cfg->build_payload_types, B200J_ERR_INVALID, "build_payload_types is NULL");
        REQUIRE(cfg->n_probe_payload == 0 || cfg->probe_payload_types, B200J_ERR_INVALID, "probe_payload_types is NULL");
        REQUIRE(cfg->layout >= 0 && cfg->layout <= 2, B200J_ERR_INVALID, "unknown layout");
        REQUIRE(cfg->radix_bits >= 0 && cfg->radix_bits <= 12, B200J_ERR_INVALID, "radix_bits must be in 0..12");
        auto ok_type = [](int t) { return t == B200J_T_UTF8 || t == 1 || t == 2 || t == 4 || t == 8 || t == 16; };
        for (int i = 0; i < cfg->n_build_payload; i++) REQUIRE(ok_type(cfg->build_payload_types[i]), B200J_ERR_UNSUPPORTED, "unsupported payload type");
        for (int i = 0; i < cfg->n_probe_payload; i++) REQUIRE(ok_type(cfg->probe_payload_types[i]), B200J_ERR_UNSUPPORTED, "unsupported payload type");
        int ndev = 0;
        if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
            cudaGetLastError();
            throw Fail{B200J_ERR_CUDA, "no CUDA device visible: this library has no CPU fallback"};
        }
        REQUIRE(cfg->device >= 0 && cfg->device < ndev, B200J_ERR_INVALID, "device ordinal out of range");
        CU(cudaSetDevice(cfg->device));
        if (const char *g = getenv("B200J_L2_FETCH_GRANULARITY")) {      // tuning knob: L2 sector promotion on random table reads
            if (cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)atoi(g)) != cudaSuccess) cudaGetLastError();
        }

        h = new b200j_handle();
        h->cfg = *cfg;
        h->key_widths.assign(cfg->key_widths, cfg->key_widths + cfg->n_keys);
        h->build_payload_types.assign(cfg->build_payload_types, cfg->build_payload_types + cfg->n_build_payload);
        h->probe_payload_types.assign(cfg->probe_payload_types, cfg->probe_payload_types + cfg->n_probe_payload);
        h->cfg.key_widths = h->key_widths.data();
        h->cfg.build_payload_types = h->build_payload_types.data();
        h->cfg.probe_payload_types = h->probe_payload_types.data();
        uint64_t k[4] = {cfg->hash_keys[0], cfg->hash_keys[1], cfg->hash_keys[2], cfg->hash_keys[3]};
        if (!(k[0] | k[1] | k[2] | k[3])) keys_with_seed(0, k);
        h->hk = HashKeys{k[0], k[1], k[2], k[3], cfg->hash_flags, 0};
        h->timing = (cfg->flags & B200J_FLAG_TIMING) != 0;
        h->composite = cfg->n_keys > 1;
        for (int i = 0; i < cfg->n_keys; i++) h->composite |= cfg->key_widths[i] == B200J_T_UTF8;     // Utf8 keys: row-hash table + equal_rows
        h->table_kw = h->composite ? 8 : cfg->key_widths[0];
        cudaDeviceProp prop{};
        CU(cudaGetDeviceProperties(&prop, cfg->device));
        h->sm_count = prop.multiProcessorCount;
        {
            int occ = 0;
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, probe_fast_kernel<8, PM_PAIRS, false, true>, PF_THREADS, 0) == cudaSuccess && occ > 0)
                h->probe_ctas_per_sm = occ;
            else
                cudaGetLastError();
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, probe_dense_kernel<8, false, true, false>, PD_THREADS, 0) == cudaSuccess && occ > 0)
                h->dense_ctas_per_sm = occ;
            else
                cudaGetLastError();
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, probe_dense_seg_kernel<8, false, false>, PD_THREADS, 0) == cudaSuccess && occ > 0)
                h->seg_ctas_per_sm = occ;
            else
                cudaGetLastError();
            if (const char *e = getenv("B200J_SEG_CTAS")) h->seg_ctas_per_sm = std::max(1, atoi(e));
            if (const char *e = getenv("B200J_NO_BLOCK_PART")) h->block_part_ok = atoi(e) == 0;
            if (const char *e = getenv("B200J_NO_NARROW")) h->narrow_ok = atoi(e) == 0;
            if (const char *e = getenv("B200J_FUSED_BUILD")) h->fused_build = atoi(e) != 0;
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, build_insert_fused_kernel<8>, TILE_THREADS, 0) == cudaSuccess && occ > 0)
                h->fused_ctas_per_sm = occ;
            else
                cudaGetLastError();
            if (const char *e = getenv("B200J_FUSED_CTAS")) h->fused_ctas_per_sm = std::max(1, atoi(e));
            if (const char *e = getenv("B200J_KEY_VIA_PROBE")) h->key_via_probe = atoi(e) != 0;
            if (const char *e = getenv("B200J_GATHER_SEG")) h->gather_seg = atoi(e) != 0;
            if (const char *e = getenv("B200J_AUTO_HINT")) h->auto_hint = atoi(e) != 0;
            if (const char *e = getenv("B200J_UTF8_LANES")) h->utf8_lanes = atoi(e);
            if (const char *e = getenv("B200J_NO_DENSE")) h->dense_ok = atoi(e) == 0;
            if (const char *e = getenv("B200J_DENSE_MISS_MAX")) h->dense_miss_max = (float)atof(e);
            if (const char *e = getenv("B200J_L1_TABLE_BYTES")) h->l1_table_bytes = (uint64_t)atoll(e);
            if (const char *e = getenv("B200J_SKEW_L1")) h->skew_l1 = atoi(e) != 0;
            if (const char *e = getenv("B200J_RADIX_LF")) h->radix_lf = std::min(0.9f, std::max(0.05f, (float)atof(e)));
            if (const char *e = getenv("B200J_THIN_LF_MIN")) h->thin_lf_min = std::min(0.5f, std::max(0.05f, (float)atof(e)));
            if (const char *e = getenv("B200J_THIN_TABLE_BYTES")) h->thin_table_bytes = std::max<uint64_t>(1, (uint64_t)atoll(e));
        }
        h->l2_bytes = prop.l2CacheSize;
        h->cache_limit = (size_t)((double)prop.totalGlobalMem * 0.4);
        if (const char *e = getenv("B200J_CACHE_LIMIT_BYTES")) h->cache_limit = (size_t)std::max<long long>(0, atoll(e));
        if (const char *e = getenv("B200J_GUARD")) h->guard = atoi(e) != 0;
        set_kernel_attributes();       // per device: opt-in shared-memory sizes of the partition kernels
        h->static_hist = static_hist_address();        // per device as well: cfg->device is current here
        if (cfg->stream) {
            h->stream = (cudaStream_t)cfg->stream;
        } else {
            CU(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
            h->own_stream = true;
        }
        CU(cudaStreamCreateWithFlags(&h->copy_in, cudaStreamNonBlocking));
        CU(cudaStreamCreateWithFlags(&h->copy_out, cudaStreamNonBlocking));
        CU(cudaEventCreate(&h->ev0));
        CU(cudaEventCreate(&h->ev1));
        // keep freed blocks in the stream-ordered pool: result buffers are recycled, not returned to the driver
        cudaMemPool_t pool;
        CU(cudaDeviceGetDefaultMemPool(&pool, cfg->device));
        uint64_t thr = UINT64_MAX;
        CU(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr));
        CU(cudaMalloc((void **)&h->d_ctr, sizeof(Counters)));
        CU(cudaHostAlloc((void **)&h->h_ctr, sizeof(Counters), cudaHostAllocDefault));
        memset(h->h_ctr, 0, sizeof(Counters));
        CU(cudaMalloc((void **)&h->d_part_rows, ((size_t)(1 << PB_MAX_BITS) + 2) * 8));
        CU(cudaHostAlloc((void **)&h->h_part_rows, ((size_t)(1 << PB_MAX_BITS) + 2) * 8, cudaHostAllocDefault));
        h->bkeys.resize(cfg->n_keys);
        for (int i = 0; i < cfg->n_keys; i++) h->bkeys[i].type = cfg->key_widths[i];
        h->bpayload.resize(cfg->n_build_payload);
        for (int i = 0; i < cfg->n_build_payload; i++) h->bpayload[i].type = cfg->build_payload_types[i];
        h->build_done.assign(cfg->n_partitions, 0);
        h->probe_done.assign(cfg->n_partitions, 0);
        h->build_cap_rows = std::max<int64_t>(cfg->expected_build_rows, 0);
        h->st.struct_size = sizeof(b200j_stats);
        if (h->build_cap_rows > 0)
            for (auto &c : h->bkeys) dev_alloc(h, c.values, (size_t)h->build_cap_rows * c.type + 16);
        *out = h;
        return B200J_OK;
    } catch (const Fail &e) {
        g_last_error = e.msg;
        delete h;
        return e.code;
    } catch (...) {
        g_last_error = "unknown error in b200j_create";
        delete h;
        return B200J_ERR_CUDA;
    }
}

int32_t b200j_destroy(b200j_handle *h) {
    if (!h) return B200J_OK;
    cudaSetDevice(h->cfg.device);
    cudaStreamSynchronize(h->stream);
    auto fr = [&](DevBuf &b) { if (b.p) big_free(h, b.p); b.p = nullptr; };
    for (auto *v : {&h->bkeys, &h->bpayload})
        for (auto &c : *v) { fr(c.values); fr(c.offsets); fr(c.validity); }
    for (auto &stg : h->stages)
        for (auto *v : {&stg.keys, &stg.payload})
            for (auto &c : *v) { fr(c.values); fr(c.offsets); fr(c.validity); }
    radix_release(h);
    fr(h->slots); fr(h->zero_slot); fr(h->rowids); fr(h->bigruns); fr(h->hot_table); fr(h->visited); fr(h->parts);
    for (auto &sc : h->slot_cols) { fr(sc.values); fr(sc.validity); }
    cudaStreamSynchronize(h->stream);
    for (auto &b : h->blocks) cudaFreeAsync(b.p, h->stream);
    h->blocks.clear();
    cudaStreamSynchronize(h->stream);
    for (auto &p : h->pinned_cache) { pinned_size(p.p, true); cudaFreeHost(p.p); }
    if (h->d_ctr) cudaFree(h->d_ctr);
    if (h->h_ctr) cudaFreeHost(h->h_ctr);
    if (h->d_part_rows) cudaFree(h->d_part_rows);
    if (h->d_reserve) cudaFree(h->d_reserve);
    for (int i = 0; i < 4; i++)
        if (h->lanes[i]) cudaStreamDestroy(h->lanes[i]);
    if (h->h_part_rows) cudaFreeHost(h->h_part_rows);
    if (h->ev0) cudaEventDestroy(h->ev0);
    if (h->ev1) cudaEventDestroy(h->ev1);
    if (h->copy_in) cudaStreamDestroy(h->copy_in);
    if (h->copy_out) cudaStreamDestroy(h->copy_out);
    if (h->own_stream) cudaStreamDestroy(h->stream);
    cudaGetLastError();
    delete h;
    return B200J_OK;
}

const char *b200j_last_error(const b200j_handle *h) { return h ? h->err.c_str() : g_last_error.c_str(); }

int32_t b200j_get_stats(b200j_handle *h, b200j_stats *out) {
    if (!h || !out) return B200J_ERR_INVALID;
    std::lock_guard<std::mutex> g(h->mu);
    const uint32_t sz = std::min<uint32_t>(out->struct_size ? out->struct_size : sizeof(b200j_stats), sizeof(b200j_stats));
    h->st.struct_size = sizeof(b200j_stats);
    memcpy(out, &h->st, sz);
    return B200J_OK;
}

int32_t b200j_reset_stats(b200j_handle *h) {
    if (!h) return B200J_ERR_INVALID;
    std::lock_guard<std::mutex> g(h->mu);
    h->st.probe_rows = h->st.result_rows = 0;
    h->st.ms_build_copy = h->st.ms_build_partition = h->st.ms_build_insert = 0;
    h->st.ms_probe_partition = h->st.ms_probe = h->st.ms_gather = 0;
    h->st.probe_launches = h->st.kernel_launches = h->st.probe_retries = 0;
    h->st.dense_probe_launches = h->st.compact_launches = h->st.seg_probe_launches = h->st.narrow_probe_launches = h->st.gather_seg_launches = h->st.gather_hint_launches = h->st.probe_order_results = 0;
    h->st.ms_slot_order = 0;
    h->st.ms_exchange_partition = 0;
    return B200J_OK;
}

int32_t b200j_reset(b200j_handle *h) {
    return guarded(h, [&] {
        CU(cudaStreamSynchronize(h->stream));
        h->n_build = 0;
        for (auto *v : {&h->bkeys, &h->bpayload})
            for (auto &c : *v) { c.has_nulls = false; c.data_bytes = 0; }
        std::fill(h->build_done.begin(), h->build_done.end(), 0);
        std::fill(h->probe_done.begin(), h->probe_done.end(), 0);
        h->n_build_done = h->n_probe_done = 0;
        h->built = false;
        h->miss_frac = 0.f;
        for (auto &stg : h->stages) {
            stg.n = 0;
            for (auto *v : {&stg.keys, &stg.payload})
                for (auto &c : *v) { c.has_nulls = false; c.data_bytes = 0; }
        }
        if (!getenv("B200J_NO_NARROW") || atoi(getenv("B200J_NO_NARROW")) == 0) h->narrow_ok = true;
        h->n_slots = 0;
        h->slot_cols_ready = false;
    });
}

int32_t b200j_synchronize(b200j_handle *h) {
    return guarded(h, [&] {
        CU(cudaStreamSynchronize(h->stream));
        REQUIRE(h->guard_violations == 0, B200J_ERR_CUDA, "B200J_GUARD: a kernel wrote outside one of the handle's buffers (see stderr)");
    });
}

int32_t b200j_trim(b200j_handle *h) {
    return guarded(h, [&] { cache_trim(h); });
}

int32_t b200j_build_push(b200j_handle *h, int32_t partition, const b200j_batch *batch) {
    return guarded(h, [&] { do_build_push(h, partition, batch); });
}
int32_t b200j_build_finish(b200j_handle *h, int32_t partition) {
    return guarded(h, [&] { do_build_finish(h, partition); });
}
int32_t b200j_probe_batch(b200j_handle *h, int32_t partition, const b200j_batch *batch, uint32_t want, b200j_result *out) {
    return guarded(h, [&] { do_probe_batch(h, partition, batch, want, out); });
}
int32_t b200j_probe_pieces(b200j_handle *h, int32_t partition, int32_t n_pieces, const void *const *piece_keys, const int64_t *piece_rows,
                           int32_t piece_key_width, uint32_t want, b200j_result *out) {
    return guarded(h, [&] { do_probe_pieces(h, partition, n_pieces, piece_keys, piece_rows, piece_key_width, want, out); });
}
int32_t b200j_apply_pairs(b200j_handle *h, int32_t partition, int64_t n_probe_rows, int32_t mem, int64_t n_pairs,
                          const uint64_t *build_idx, const uint32_t *probe_idx, uint32_t want, b200j_result *out) {
    return guarded(h, [&] { do_apply_pairs(h, partition, n_probe_rows, mem, n_pairs, build_idx, probe_idx, want, out); });
}
int32_t b200j_probe_finish(b200j_handle *h, int32_t partition, uint32_t want, b200j_result *out) {
    return guarded(h, [&] { do_probe_finish(h, partition, want, out); });
}
int32_t b200j_result_release(b200j_handle *h, b200j_result *r) {
    return guarded(h, [&] {
        if (!r) return;
        release_owner(h, (ResultOwner *)r->_owner);
        r->_owner = nullptr;
        r->build_idx = nullptr;
        r->probe_idx = nullptr;
        r->n = 0;
    });
}
int32_t b200j_resolve_origin(b200j_handle *h, int64_t n, const void *idx, int32_t idx_width, int32_t n_pieces, const int64_t *piece_rows,
                             const uint32_t *const *piece_rowid, const uint32_t *piece_rowid_base, uint32_t *out_piece, uint32_t *out_row) {
    return guarded(h, [&] { do_resolve_origin(h, n, idx, idx_width, n_pieces, piece_rows, piece_rowid, piece_rowid_base, out_piece, out_row); });
}

int32_t b200j_gather_pieces(b200j_handle *h, int64_t n, const uint32_t *piece, const uint32_t *row, int32_t n_pieces, const void *const *col_values,
                            const uint8_t *const *col_validity, int32_t width, void *out_values, uint32_t *out_validity) {
    return guarded(h, [&] { do_gather_pieces(h, n, piece, row, n_pieces, col_values, col_validity, width, out_values, out_validity); });
}

int32_t b200j_gather_pieces_utf8_offsets(b200j_handle *h, int64_t n, const uint32_t *piece, const uint32_t *row, int32_t n_pieces,
                                         const int32_t *const *col_offsets, const uint8_t *const *col_validity, int32_t *out_offsets,
                                         uint32_t *out_validity, int64_t *total_bytes) {
    return guarded(h, [&] { do_gather_pieces_utf8_offsets(h, n, piece, row, n_pieces, col_offsets, col_validity, out_offsets, out_validity, total_bytes); });
}

int32_t b200j_gather_pieces_utf8_data(b200j_handle *h, int64_t n, const uint32_t *piece, const uint32_t *row, int32_t n_pieces,
                                      const int32_t *const *col_offsets, const uint8_t *const *col_data, const int32_t *out_offsets, uint8_t *out_data) {
    return guarded(h, [&] { do_gather_pieces_utf8_data(h, n, piece, row, n_pieces, col_offsets, col_data, out_offsets, out_data); });
}

int32_t b200j_utf8_rows_equal(b200j_handle *h, int64_t n, const int32_t *off_a, const uint8_t *data_a, const uint32_t *valid_a, const int32_t *off_b,
                              const uint8_t *data_b, const uint32_t *valid_b, uint8_t *out) {
    return guarded(h, [&] { do_utf8_rows_equal(h, n, off_a, data_a, valid_a, off_b, data_b, valid_b, out); });
}

int32_t b200j_gather(b200j_handle *h, const b200j_result *pairs, const b200j_batch *probe_batch, uint32_t want, b200j_gathered *out) {
    return guarded(h, [&] { do_gather(h, pairs, probe_batch, want, out); });
}
int32_t b200j_gathered_release(b200j_handle *h, b200j_gathered *g) {
    return guarded(h, [&] { do_gathered_release(h, g); });
}
int32_t b200j_probe_push(b200j_handle *h, int32_t partition, const b200j_batch *batch) {
    return guarded(h, [&] { do_probe_push(h, partition, batch); });
}
int64_t b200j_staged_rows(b200j_handle *h, int32_t partition) {
    if (!h) return -1;
    std::lock_guard<std::mutex> g(h->mu);
    return (partition >= 0 && partition < (int32_t)h->stages.size()) ? h->stages[partition].n : 0;
}
int32_t b200j_probe_flush(b200j_handle *h, int32_t partition, uint32_t want, b200j_result *out) {
    return guarded(h, [&] { do_probe_flush(h, partition, want, out); });
}
int32_t b200j_gather_staged(b200j_handle *h, int32_t partition, const b200j_result *pairs, uint32_t want, b200j_gathered *out) {
    return guarded(h, [&] { do_gather_staged(h, partition, pairs, want, out); });
}
int32_t b200j_hash_keys(b200j_handle *h, const b200j_batch *batch, uint64_t *out) {
    return guarded(h, [&] { do_hash_keys(h, batch, out); });
}

int32_t b200j_shuffle_count_static(b200j_handle *h, const void *keys, const uint8_t *validity, int64_t bit_offset, int64_t n_rows,
                                   int32_t n_dest, int64_t *counts) {
    return guarded(h, [&] { do_shuffle_count_static(h, keys, validity, bit_offset, n_rows, n_dest, counts); });
}

int32_t b200j_shuffle_count(b200j_handle *h, const b200j_batch *batch, int32_t n_dest, int64_t *counts) {
    return guarded(h, [&] { do_shuffle_count(h, batch, n_dest, counts); });
}
int32_t b200j_shuffle_set_hot(b200j_handle *h, const uint64_t *keys, int32_t n) {
    return guarded(h, [&] { do_shuffle_set_hot(h, keys, n); });
}
int32_t b200j_shuffle_scatter(b200j_handle *h, int32_t n_dest, void *const *dest_keys, uint32_t *const *dest_rowid, const int64_t *dest_offset,
                              uint32_t rowid_base) {
    return guarded(h, [&] { do_shuffle_scatter(h, n_dest, dest_keys, dest_rowid, dest_offset, rowid_base); });
}
int32_t b200j_shuffle_local_scatter(b200j_handle *h, const b200j_batch *batch, int32_t n_dest, void *send_keys, uint32_t *send_rowid,
                                    int64_t region_rows, uint32_t rowid_base, int32_t key_out_width, int64_t *counts) {
    return guarded(h, [&] { do_shuffle_local_scatter(h, batch, n_dest, send_keys, send_rowid, region_rows, rowid_base, key_out_width, counts); });
}
int32_t b200j_copy_async(b200j_handle *h, void *dst, const void *src, size_t bytes, int32_t lane) {
    return guarded(h, [&] {
        REQUIRE(lane >= 0 && lane < 4, B200J_ERR_INVALID, "copy lane must be in 0..3");
        if (bytes == 0) return;
        REQUIRE(dst && src, B200J_ERR_INVALID, "b200j_copy_async: NULL pointer");
        if (!h->lanes[lane]) CU(cudaStreamCreateWithFlags(&h->lanes[lane], cudaStreamNonBlocking));
        cudaEvent_t ev = nullptr;                       // the copy starts after what the handle's stream holds now
        CU(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        cudaEventRecord(ev, h->stream);
        cudaStreamWaitEvent(h->lanes[lane], ev, 0);
        cudaEventDestroy(ev);
        CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, h->lanes[lane]));
    });
}
int32_t b200j_copy_wait(b200j_handle *h) {
    return guarded(h, [&] {
        for (int i = 0; i < 4; i++)
            if (h->lanes[i]) CU(cudaStreamSynchronize(h->lanes[i]));
    });
}
int32_t b200j_ipc_alloc(b200j_handle *h, size_t bytes, void **ptr, uint8_t *handle_out) {
    return guarded(h, [&] {
        REQUIRE(ptr && handle_out, B200J_ERR_INVALID, "b200j_ipc_alloc: NULL argument");
        static_assert(sizeof(cudaIpcMemHandle_t) == B200J_IPC_HANDLE_BYTES, "cudaIpcMemHandle_t is 64 bytes");
        *ptr = nullptr;
        CU(cudaMalloc(ptr, std::max<size_t>(bytes, 256)));      // legacy IPC needs cudaMalloc memory, not the stream-ordered pool
        cudaIpcMemHandle_t ih;
        cudaError_t e = cudaIpcGetMemHandle(&ih, *ptr);
        if (e != cudaSuccess) {
            cudaFree(*ptr);
            *ptr = nullptr;
            CU(e);
        }
        memcpy(handle_out, &ih, sizeof(ih));
    });
}
int32_t b200j_ipc_open(b200j_handle *h, const uint8_t *handle, void **peer_ptr) {
    return guarded(h, [&] {
        REQUIRE(handle && peer_ptr, B200J_ERR_INVALID, "b200j_ipc_open: NULL argument");
        cudaIpcMemHandle_t ih;
        memcpy(&ih, handle, sizeof(ih));
        CU(cudaIpcOpenMemHandle(peer_ptr, ih, cudaIpcMemLazyEnablePeerAccess));
    });
}
int32_t b200j_ipc_close(b200j_handle *h, void *peer_ptr) {
    return guarded(h, [&] { if (peer_ptr) CU(cudaIpcCloseMemHandle(peer_ptr)); });
}
int32_t b200j_ipc_free(b200j_handle *h, void *ptr) {
    return guarded(h, [&] { if (ptr) CU(cudaFree(ptr)); });
}

int32_t b200j_host_alloc(size_t bytes, void **out) {
    if (!out) return B200J_ERR_INVALID;
    *out = nullptr;
    cudaError_t e = cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault);
    if (e != cudaSuccess) {
        cudaGetLastError();
        g_last_error = std::string("cudaHostAlloc: ") + cudaGetErrorString(e);
        return e == cudaErrorMemoryAllocation ? B200J_ERR_OOM : B200J_ERR_CUDA;
    }
    return B200J_OK;
}
int32_t b200j_host_free(void *p) {
    if (p && cudaFreeHost(p) != cudaSuccess) {
        cudaGetLastError();
        return B200J_ERR_CUDA;
    }
    return B200J_OK;
}

}  // extern "C"
