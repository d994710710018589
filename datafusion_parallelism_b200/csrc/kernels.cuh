// kernels.cuh — sm_100a device code of the B200 hash join (no tensor cores: nothing here is a contraction;
// every kernel is HBM/L2-bound integer work, so the rules that matter are coalesced 128-bit loads,
// warp-aggregated atomics, and keeping the random-access working set inside the 126 MB L2).
//
// What each kernel replaces in the reference (jamesfer/datafusion-parallelism @ 3e184de):
//   hash_one / hash_keys_kernel      calculate_hash, src/shared/shared.rs:11-16 (create_hashes over ahash)
//   build_insert_kernel              JoinStateInstance::add, version10/parallel_join_execution_state.rs:91-133
//                                    + Inner::insert_atomically, version10/new_map_3/fixed_table.rs:560-672
//   csr_scan_kernel/csr_fill_kernel  the overflow chain of version10 (parallel_join_execution_state.rs:101-106),
//                                    re-laid out as contiguous row-id runs so long duplicate chains are copied
//                                    warp-wide instead of pointer-chased
//   probe_kernel                     get_matching_indices[_with_probe] (shared.rs:29-81) + Version10Lookup::get_iter
//                                    (lookup_implementation_3.rs:31-59) + equal_rows_arr (datafusion_private.rs:40-80,
//                                    fused: slots hold the key VALUE, so a slot match is a key match) +
//                                    append_right_indices / get_semi_indices / get_anti_indices
//                                    (datafusion_private.rs:85-185) + ConcurrentBitSet::set_ones (concurrent_bit_set.rs:28-52)
//   bitmap_select_kernel             get_unset_indices_array / get_set_indices_array (concurrent_bit_set.rs:93-131)
#pragma once
#include <cstdint>
#include <type_traits>
#include <cuda_runtime.h>

namespace b200j {

// ------------------------------------------------------------------------------------------------
// ahash 0.8.11 fallback arithmetic (bit-identical to oracle/join_oracle.c)
// ------------------------------------------------------------------------------------------------
constexpr uint64_t AHASH_MULTIPLE = 6364136223846793005ULL;

struct HashKeys {
    uint64_t k0, k1, k2, k3;
    int32_t  flags;
    int32_t  _pad;
};

// 64x64 -> 128 bit product folded to 64 bits (low ^ high). Written as mul.lo ^ mul.hi: ptxas turns the pair into
// four IMAD.WIDE.U32 plus three carry instructions (9 SASS instructions; a hand-written limb version compiled to 21).
__host__ __device__ __forceinline__ uint64_t folded_multiply(uint64_t s, uint64_t by) {
#ifdef __CUDA_ARCH__
    return (s * by) ^ __umul64hi(s, by);
#else
    unsigned __int128 r = (unsigned __int128)s * (unsigned __int128)by;
    return (uint64_t)r ^ (uint64_t)(r >> 64);
#endif
}

__host__ __device__ __forceinline__ uint64_t rotl64(uint64_t x, unsigned r) {
    r &= 63u;
    return r ? (x << r) | (x >> (64u - r)) : x;
}

// hash_one of a <=64-bit integer (32-bit keys arrive zero-extended). flags bit0: stable-toolchain variant.
__host__ __device__ __forceinline__ uint64_t hash_one(const HashKeys &k, uint64_t v) {
    const uint64_t buffer = folded_multiply(v ^ k.k1, AHASH_MULTIPLE);
    const uint64_t h = folded_multiply(buffer, k.k0);
    return (k.flags & 1) ? rotl64(h, (unsigned)(buffer & 63u)) : h;
}

// hash_one of a Utf8 value: std's `Hash for str` (write(bytes), write_u8(0xff)) over ahash's fallback AHasher
// {buffer: k1, pad: k0, extra: [k2, k3]} -- restated in oracle/join_oracle.c orc_hash_bytes, which this must equal
// bit for bit. Unaligned little-endian reads are assembled from bytes (strings start at any offset).
__device__ __forceinline__ uint64_t load_le(const uint8_t *p, int n) {
    uint64_t v = 0;
    for (int i = 0; i < n; i++) v |= (uint64_t)p[i] << (8 * i);
    return v;
}
__device__ __forceinline__ uint64_t ah_large_update(const HashKeys &k, uint64_t buffer, uint64_t a, uint64_t b) {
    return rotl64((buffer + k.k0) ^ folded_multiply(a ^ k.k2, b ^ k.k3), 23u);
}
__device__ inline uint64_t hash_bytes(const HashKeys &k, const uint8_t *d, uint64_t len) {
    uint64_t buffer = (k.k1 + len) * AHASH_MULTIPLE;
    if (len > 8) {
        if (len > 16) {
            buffer = ah_large_update(k, buffer, load_le(d + len - 16, 8), load_le(d + len - 8, 8));
            for (uint64_t off = 0; len - off > 16; off += 16)
                buffer = ah_large_update(k, buffer, load_le(d + off, 8), load_le(d + off + 8, 8));
        } else {
            buffer = ah_large_update(k, buffer, load_le(d, 8), load_le(d + len - 8, 8));
        }
    } else {
        uint64_t a = 0, b = 0;
        if (len >= 4) { a = load_le(d, 4); b = load_le(d + len - 4, 4); }
        else if (len >= 2) { a = load_le(d, 2); b = d[len - 1]; }
        else if (len == 1) { a = d[0]; b = d[0]; }
        buffer = ah_large_update(k, buffer, a, b);
    }
    buffer = folded_multiply(0xffULL ^ buffer, AHASH_MULTIPLE);
    return rotl64(folded_multiply(buffer, k.k0), (unsigned)(buffer & 63u));
}

// datafusion-common hash_utils::combine_hashes
__host__ __device__ __forceinline__ uint64_t combine_hashes(uint64_t l, uint64_t r) {
    return ((uint64_t)(17 * 37) + l) * 37ULL + r;
}

// ------------------------------------------------------------------------------------------------
// data descriptors
// ------------------------------------------------------------------------------------------------
constexpr int MAX_KEYS = 8;

struct KeyCol {
    const void    *values;     // fixed-width values, or the Utf8 data bytes
    const uint8_t *validity;   // LSB-first bitmap, may be null
    int64_t        bit_offset;
    int32_t        width;      // 4 or 8; 0 = Utf8
    int32_t        off_width;  // Utf8: 4 (Arrow int32 offsets of a batch) or 8 (the build store's int64 offsets)
    const void    *offsets;    // Utf8: n + 1 offsets into `values`
};

struct KeyCols {
    KeyCol  c[MAX_KEYS];
    int32_t n;
    int32_t _pad;
};

__device__ __forceinline__ bool bit_valid(const uint8_t *bm, int64_t i) {
    return bm == nullptr || ((bm[i >> 3] >> (i & 7)) & 1);
}

__device__ __forceinline__ int64_t utf8_offset(const KeyCol &kc, int64_t i) {
    return kc.off_width == 8 ? reinterpret_cast<const int64_t *>(kc.offsets)[i] : (int64_t)reinterpret_cast<const int32_t *>(kc.offsets)[i];
}

// hash_one of row i of one key column (the value is known to be non-NULL)
__device__ __forceinline__ uint64_t hash_col_value(const HashKeys &hk, const KeyCol &kc, int64_t i) {
    if (kc.width == 0) {
        const int64_t o0 = utf8_offset(kc, i), o1 = utf8_offset(kc, i + 1);
        return hash_bytes(hk, reinterpret_cast<const uint8_t *>(kc.values) + o0, (uint64_t)(o1 - o0));
    }
    const uint64_t v = kc.width == 8 ? reinterpret_cast<const unsigned long long *>(kc.values)[i]
                                     : (uint64_t)reinterpret_cast<const uint32_t *>(kc.values)[i];
    return hash_one(hk, v);
}

// Table slot: 16 B, two per 32 B sector. key == 0 marks an empty slot (the key value 0 lives in a side slot).
//   count == 1 : `row` is the build row
//   count >= 2 : `row` is the start of a run of `count` build rows in rowids[]
struct __align__(16) Slot {
    unsigned long long key;
    uint32_t           row;
    uint32_t           count;
};

// Top bit of the EVEN slot's count: "a key whose home bucket this is was placed further on" (set by the insert that leaves its
// full home bucket). A lookup that does not find its key in a full home bucket whose bit is clear has proven a miss: every key
// homed there is still there. Only home buckets are asked; count readers mask the bit.
constexpr uint32_t SLOT_OVF   = 0x80000000u;
constexpr uint32_t COUNT_MASK = 0x7fffffffu;
__device__ __forceinline__ bool bucket_overflowed(unsigned long long v0) { return (v0 >> 63) != 0ULL; }

struct Counters {
    // reset before every probe launch (24 contiguous bytes)
    unsigned long long out_count;     // rows produced (may exceed capacity: then nothing past capacity is written)
    unsigned long long build_null_out;
    unsigned int       next_tile;     // dynamic tile scheduler
    unsigned int       big_count;     // duplicate runs deferred to expand_runs_kernel (may exceed the list's capacity)
    // build
    unsigned long long distinct;
    unsigned long long dup_rows;
    unsigned long long multi_slots;
    unsigned long long null_keys;
    unsigned long long csr_cursor;
};
constexpr size_t COUNTERS_LAUNCH_BYTES = 24;

// One radix partition's table: buckets [base, base + nb) of the slot array (bucket = 2 slots = one 32 B sector).
struct __align__(8) PartEnt {
    uint32_t base;
    uint32_t nb;
};

struct TableView {
    Slot           *slots;
    Slot           *zero_slot;     // the bucket of key value 0
    uint32_t       *rowids;        // runs of duplicate-key rows
    const PartEnt  *parts;         // radix layout: partition p = hash >> (64 - radix_bits)
    uint32_t        nb;            // direct layout: number of buckets
    int32_t         radix_bits;
    uint32_t        l1;            // probe loads may allocate in L1 (small direct table)
    uint32_t        _pad;
};

// hp = top 32 bits of the row hash. Table sizes are not powers of two (the load factor is what the caller asked
// for, not what rounding gives): the home bucket is mulhi(low 32 hash bits, nb). Tables never exceed 2^32 buckets.
template <bool RADIX>
__device__ __forceinline__ void locate(const TableView &tv, uint32_t hp, uint32_t &base, uint32_t &nb) {
    if (RADIX) {
        const uint2 e = __ldg(reinterpret_cast<const uint2 *>(tv.parts + (hp >> (32 - tv.radix_bits))));
        base = e.x;
        nb = e.y;
    } else {
        base = 0;
        nb = tv.nb;
    }
}

__device__ __forceinline__ uint32_t home_bucket(uint64_t h, uint32_t nb) { return __umulhi((uint32_t)h, nb); }
__device__ __forceinline__ uint32_t next_bucket(uint32_t b, uint32_t nb) { return b + 1u == nb ? 0u : b + 1u; }
__device__ __forceinline__ Slot *bucket_ptr(const TableView &tv, uint32_t base, uint32_t b) {
    return tv.slots + 2ull * ((unsigned long long)base + b);
}

__device__ __forceinline__ Slot load_slot(const Slot *p) {
    // 16 B load that bypasses L1 (random access: L1 only pollutes) and reads through L2.
    const ulonglong2 v = __ldcg(reinterpret_cast<const ulonglong2 *>(p));
    Slot s;
    s.key = v.x;
    s.row = (uint32_t)v.y;
    s.count = (uint32_t)(v.y >> 32) & COUNT_MASK;
    return s;
}

// One 32 B sector = one bucket of two slots, fetched with a single 256-bit load (LDG.E.256, sm_100+) that
// bypasses L1 and asks L2 to keep the line (evict_last): the table is the only data with reuse, everything
// else in the probe is a stream marked evict-first.
// Linear probing always starts at the even slot of the home bucket, so a probe step is exactly one sector.
// l1 (TableView::l1, warp-uniform): a table of a few hundred KB at most (dimension tables: nation, region ...) is read
// through L1 instead -- every SM otherwise hammers the same handful of L2 sectors (measured: 25 build rows probed at
// 30 G rows/s, slower than a 30 MB table). The table is read-only while probes run, so L1 cannot go stale.
__device__ __forceinline__ void load_bucket(const Slot *p, unsigned long long &k0, unsigned long long &v0,
                                            unsigned long long &k1, unsigned long long &v1, bool l1) {
    if (l1) asm volatile("ld.global.L1::evict_last.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(k0), "=l"(v0), "=l"(k1), "=l"(v1) : "l"(p));
    else asm volatile("ld.global.L1::no_allocate.L2::evict_last.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(k0), "=l"(v0), "=l"(k1), "=l"(v1) : "l"(p));
}

// ------------------------------------------------------------------------------------------------
// vectorised key loads: 4 keys per thread per tile, 128-bit loads wherever the address allows
// ------------------------------------------------------------------------------------------------
constexpr int TILE_THREADS = 256;
constexpr int TILE_ITEMS   = 4;
constexpr int TILE_ROWS    = TILE_THREADS * TILE_ITEMS;

// `keys` is the caller's pointer; `mis` = number of keys by which it is misaligned from 16 B. Rows are
// addressed in the aligned "virtual" index space v = row + mis, so every full vector load is 16 B aligned.
// A tile is THREADS * ITEMS consecutive rows; each thread takes ITEMS * KW / 16 vectors, THREADS vectors apart,
// so a warp's loads are 512 contiguous bytes.
template <int KW, int THREADS, int ITEMS>
__device__ __forceinline__ void load_keys_tile(const void *keys, int mis, int64_t n, int64_t tile, int t,
                                               uint64_t (&key)[ITEMS], int64_t (&row)[ITEMS]) {
    const int64_t vbase = tile * (int64_t)(THREADS * ITEMS);
    const int64_t vend = n + mis;
    if (KW == 8) {
        const unsigned long long *k64 = reinterpret_cast<const unsigned long long *>(keys) - mis;
#pragma unroll
        for (int v = 0; v < ITEMS / 2; v++) {
            const int64_t vi = vbase + ((int64_t)v * THREADS + t) * 2;
            row[2 * v] = vi - mis;
            row[2 * v + 1] = vi + 1 - mis;
            if (vi >= mis && vi + 1 < vend) {
                const ulonglong2 x = __ldcs(reinterpret_cast<const ulonglong2 *>(k64 + vi));
                key[2 * v] = x.x;
                key[2 * v + 1] = x.y;
            } else {
                key[2 * v] = (vi >= mis && vi < vend) ? __ldcs(k64 + vi) : 0ULL;
                key[2 * v + 1] = (vi + 1 >= mis && vi + 1 < vend) ? __ldcs(k64 + vi + 1) : 0ULL;
            }
        }
    } else {
        const uint32_t *k32 = reinterpret_cast<const uint32_t *>(keys) - mis;
#pragma unroll
        for (int v = 0; v < ITEMS / 4; v++) {
            const int64_t vi = vbase + ((int64_t)v * THREADS + t) * 4;
#pragma unroll
            for (int q = 0; q < 4; q++) row[4 * v + q] = vi + q - mis;
            if (vi >= mis && vi + 3 < vend) {
                const uint4 x = __ldcs(reinterpret_cast<const uint4 *>(k32 + vi));
                key[4 * v] = x.x; key[4 * v + 1] = x.y; key[4 * v + 2] = x.z; key[4 * v + 3] = x.w;
            } else {
#pragma unroll
                for (int q = 0; q < 4; q++)
                    key[4 * v + q] = (vi + q >= mis && vi + q < vend) ? (uint64_t)__ldcs(k32 + vi + q) : 0ULL;
            }
        }
    }
}

template <int KW>
__device__ __forceinline__ void load_tile_keys(const void *keys, int mis, int64_t n, int64_t tile, int t,
                                               uint64_t (&key)[TILE_ITEMS], int64_t (&row)[TILE_ITEMS]) {
    load_keys_tile<KW, TILE_THREADS, TILE_ITEMS>(keys, mis, n, tile, t, key, row);
}

__device__ __forceinline__ unsigned long long warp_sum_u32(uint32_t v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ------------------------------------------------------------------------------------------------
// a1: create_hashes
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) hash_keys_kernel(KeyCols cols, int64_t n, HashKeys hk, uint64_t *__restrict__ out) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        uint64_t h = 0;
        for (int c = 0; c < cols.n; c++) {
            const KeyCol &kc = cols.c[c];
            if (!bit_valid(kc.validity, kc.bit_offset + i)) continue;     // NULL leaves the running hash unchanged
            const uint64_t hv = hash_col_value(hk, kc, i);
            h = (c == 0) ? hv : combine_hashes(hv, h);
        }
        out[i] = h;
    }
}

// Composite join keys (n_keys > 1): the table is keyed on the 64-bit ROW HASH, exactly like the reference's map
// (version10 stores hashes, new_map_3.rs:250-303), and key equality is checked afterwards on the candidate pairs
// (equal_rows_arr, datafusion_private.rs:40-80). This kernel produces the row hashes (create_hashes over all key
// columns, combine_hashes from the second column on) and the "joinable" bitmap: a row with a NULL in ANY key column
// can never match (null_equals_null = false, inner.rs:105-107), so it is kept out of the table / the lookup.
__global__ void __launch_bounds__(256) composite_hash_kernel(KeyCols cols, int64_t n, HashKeys hk, uint64_t *__restrict__ out,
                                                             uint32_t *__restrict__ joinable) {
    const int lane = threadIdx.x & 31;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x - lane; i0 < n; i0 += stride) {
        const int64_t i = i0 + lane;
        bool ok = i < n;
        uint64_t h = 0;
        if (i < n) {
            for (int c = 0; c < cols.n; c++) {
                const KeyCol &kc = cols.c[c];
                if (!bit_valid(kc.validity, kc.bit_offset + i)) { ok = false; continue; }
                const uint64_t hv = hash_col_value(hk, kc, i);
                h = (c == 0) ? hv : combine_hashes(hv, h);
            }
            if (hk.flags & 2) h &= 0x3ffULL;       // B200J_HASH_TEST_COLLIDE: 1024 hash values, so equal_rows has work to do
            out[i] = h;
        }
        const uint32_t w = __ballot_sync(0xffffffffu, ok);
        if (lane == 0) joinable[i0 >> 5] = w;
    }
}

// equal_rows_arr on candidate pairs: a pair survives when every key column is non-NULL and equal on both sides.
// Failed pairs get build index ~0 (compacted away by compact_pairs_kernel); their number goes to *n_failed.
__global__ void __launch_bounds__(256) equal_rows_kernel(KeyCols build, KeyCols probe, uint64_t *__restrict__ build_idx,
                                                         const uint32_t *__restrict__ probe_idx, int64_t n, unsigned long long *n_failed) {
    uint32_t failed = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t b = (int64_t)build_idx[i], p = (int64_t)probe_idx[i];
        bool eq = true;
        for (int c = 0; c < build.n && eq; c++) {
            const KeyCol &bc = build.c[c], &pc = probe.c[c];
            if (!bit_valid(bc.validity, bc.bit_offset + b) || !bit_valid(pc.validity, pc.bit_offset + p)) { eq = false; break; }
            if (bc.width == 0) {
                const int64_t bo = utf8_offset(bc, b), bl = utf8_offset(bc, b + 1) - bo;
                const int64_t po = utf8_offset(pc, p), pl = utf8_offset(pc, p + 1) - po;
                eq = bl == pl;
                const uint8_t *x = reinterpret_cast<const uint8_t *>(bc.values) + bo, *y = reinterpret_cast<const uint8_t *>(pc.values) + po;
                for (int64_t j = 0; j < bl && eq; j++) eq = x[j] == y[j];
            } else if (bc.width == 8) eq = reinterpret_cast<const unsigned long long *>(bc.values)[b] == reinterpret_cast<const unsigned long long *>(pc.values)[p];
            else eq = reinterpret_cast<const uint32_t *>(bc.values)[b] == reinterpret_cast<const uint32_t *>(pc.values)[p];
        }
        if (!eq) { build_idx[i] = ~0ULL; failed++; }
    }
    const unsigned long long f = __reduce_add_sync(0xffffffffu, failed);
    if ((threadIdx.x & 31) == 0 && f) atomicAdd(n_failed, f);
}

// ------------------------------------------------------------------------------------------------
// a2: build — open addressing, atomicCAS claim on the key word, atomicAdd on the slot's count
// ------------------------------------------------------------------------------------------------
struct BuildArgs {
    const void     *keys;
    const uint8_t  *validity;     // build store bitmap (bit 0 = row 0) or null
    const uint32_t *rowid;        // radix layout: explicit row ids of the partitioned keys; null = identity
    int64_t         n;
    int32_t         mis;
    int32_t         _pad;
    TableView       tv;
    HashKeys        hk;
    Counters       *ctr;
    uint32_t       *slot_of;      // [n] slot index every row went to (0xffffffff = the zero-key side slot), for csr_fill; may be null
    uint32_t        dup_bit;      // 1: slot indices fit 31 bits and the insert sets bit 31 of slot_of[] for the rows that found their key already
                                  //    present -- the only rows csr_fill has to place (it then skips the others without looking at their slot)
    uint32_t        _pad2;
};
constexpr uint32_t SLOT_OF_DUP = 0x80000000u;

// 128-bit compare-and-swap of a whole slot: {0, 0} -> {key, row | count 1}. One atomic claims the slot, stores the row id
// and sets the count; returns the key word found (0 = claimed). The insert kernel is bound by L2 sector operations (ncu: 11
// per row, 1.14 G per 100 M rows against the chip's ~197 per clock): claim + count + row store in ONE operation instead of
// three took the kernel from 4.5 to 3.7 ms.
__device__ __forceinline__ unsigned long long claim_slot128(Slot *s, unsigned long long key, uint32_t row) {
    unsigned long long ok, ov;
    const unsigned long long val = (unsigned long long)row | (1ULL << 32);
    asm volatile(
        "{\n\t.reg .b128 cmp, swp, old;\n\t"
        "mov.b128 cmp, {%3, %3};\n\t"
        "mov.b128 swp, {%4, %5};\n\t"
        "atom.global.relaxed.gpu.cas.b128 old, [%2], cmp, swp;\n\t"
        "mov.b128 {%0, %1}, old;\n\t}"
        : "=l"(ok), "=l"(ov)
        : "l"(s), "l"(0ULL), "l"(key), "l"(val)
        : "memory");
    return ov == ~0ULL ? ok : ok;          // (ov is the value word that came back with the key: unused)
}

template <bool RADIX, bool CLAIM128 = false>
__device__ __forceinline__ Slot *find_or_claim(const TableView &tv, const HashKeys &hk, uint64_t key, uint32_t &distinct, uint32_t row = 0) {
    if (key == 0) return tv.zero_slot;
    const uint64_t h = hash_one(hk, key);
    uint32_t base, nb;
    locate<RADIX>(tv, (uint32_t)(h >> 32), base, nb);
    uint32_t b = home_bucket(h, nb);
    for (bool home = true;; home = false) {
        Slot *s = bucket_ptr(tv, base, b);      // slots of a bucket fill in order: even first
        // look first, then claim: claiming without the pre-read was measured slower (3.98 vs 3.70 ms / 100 M rows) and lets
        // the rows of a heavy hitter hammer its slot with compare-and-swaps (Zipf build 12.8 -> 33 ms). The look is ONE 32 B
        // load of the whole bucket (one L2 request, where two 8 B key loads were two), and it brings the overflow bit along.
        unsigned long long k[2], v0, v1;
        asm volatile("ld.global.cg.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(k[0]), "=l"(v0), "=l"(k[1]), "=l"(v1) : "l"(s) : "memory");
#pragma unroll
        for (int j = 0; j < 2; j++, s++) {
            unsigned long long cur = k[j];
            if (cur == 0ULL) {
#if defined(B200J_INS_TEST) && B200J_INS_TEST == 1
                // TIMING EXPERIMENT ONLY (racy): plain 16-byte store instead of the compare-and-swap
                *reinterpret_cast<ulonglong2 *>(s) = make_ulonglong2((unsigned long long)key, (unsigned long long)row | (1ULL << 32));
                cur = 0ULL;
#elif defined(B200J_INS_TEST) && B200J_INS_TEST == 2
                cur = atomicCAS(&s->key, 0ULL, (unsigned long long)key);          // 64-bit claim, row / count stored separately
                if (cur == 0ULL) { s->row = row; s->count = 1u; }
#else
                cur = CLAIM128 ? claim_slot128(s, key, row) : atomicCAS(&s->key, 0ULL, (unsigned long long)key);
#endif
                if (cur == 0ULL) { distinct++; return s; }
            }
            if (cur == key) return s;
        }
        // leaving the full home bucket: tell the probes that it overflowed (v0 is from before our claims: a stale clear bit
        // only costs a redundant atomicOr)
        if (home && !bucket_overflowed(v0)) atomicOr(&bucket_ptr(tv, base, b)->count, SLOT_OVF);
        b = next_bucket(b, nb);
    }
}

template <int KW, bool RADIX>
__global__ void __launch_bounds__(TILE_THREADS) build_insert_kernel(BuildArgs a) {
    constexpr bool CLAIM128 = true;
    const int t = threadIdx.x;
    uint32_t distinct = 0, dups = 0, multi = 0, nulls = 0;
    const int64_t n_tiles = (a.n + a.mis + TILE_ROWS - 1) / TILE_ROWS;
    __shared__ unsigned int s_tile;
    for (;;) {
        // dynamic tile order (see probe_kernel): keeps the running CTAs inside one or two radix partitions
        __syncthreads();
        if (t == 0) s_tile = atomicAdd(&a.ctr->next_tile, 1u);
        __syncthreads();
        const int64_t tile = s_tile;
        if (tile >= n_tiles) break;
        uint64_t key[TILE_ITEMS];
        int64_t row[TILE_ITEMS];
        load_tile_keys<KW>(a.keys, a.mis, a.n, tile, t, key, row);
#pragma unroll
        for (int q = 0; q < TILE_ITEMS; q++) {
            if (row[q] < 0 || row[q] >= a.n) continue;
            if (!bit_valid(a.validity, row[q])) { nulls++; continue; }
            const uint32_t r = a.rowid ? __ldcs(a.rowid + row[q]) : (uint32_t)row[q];
            const uint32_t d0 = distinct;
            Slot *s = find_or_claim<RADIX, CLAIM128>(a.tv, a.hk, key[q], distinct, r);
            if (a.slot_of) a.slot_of[row[q]] = key[q] ? (uint32_t)(s - a.tv.slots) : 0xffffffffu;
            if (CLAIM128 && distinct != d0) continue;       // the claim wrote key, row id and count 1 in one atomic
            // Rows that found their key already present may be many rows of ONE key (a heavy hitter fills most of its radix
            // partition's tiles): lanes on the same slot add up first and send one atomic. Unique keys never take this path.
            uint32_t pos;
            const unsigned am = __activemask();
            if (__any_sync(am, distinct == d0)) {
                const unsigned peers = __match_any_sync(am, (unsigned long long)s);
                const int leader = __ffs(peers) - 1;
                uint32_t base = 0;
                if ((t & 31) == leader) base = atomicAdd(&s->count, (uint32_t)__popc(peers)) & COUNT_MASK;
                base = __shfl_sync(peers, base, leader);
                pos = base + __popc(peers & ((1u << (t & 31)) - 1u));
            } else {
                pos = atomicAdd(&s->count, 1u) & COUNT_MASK;
            }
            if (pos == 0) {
                s->row = r;
                if (key[q] == 0) distinct++;
            } else {
                dups++;
                if (pos == 1) multi++;
                if (a.dup_bit && a.slot_of && key[q]) a.slot_of[row[q]] = (uint32_t)(s - a.tv.slots) | SLOT_OF_DUP;
            }
        }
    }
    const unsigned long long d = warp_sum_u32(distinct), u = warp_sum_u32(dups), m = warp_sum_u32(multi), z = warp_sum_u32(nulls);
    if ((t & 31) == 0) {
        if (d) atomicAdd(&a.ctr->distinct, d);
        if (u) atomicAdd(&a.ctr->dup_rows, u);
        if (m) atomicAdd(&a.ctr->multi_slots, m);
        if (z) atomicAdd(&a.ctr->null_keys, z);
    }
}

// ------------------------------------------------------------------------------------------------
// Radix layout, round 2: table clear and insert in ONE persistent kernel. The separate cudaMemset of the whole slot array
// (3.8 GB at C2) wrote zeros to HBM that the insert kernel then pulled back into L2 line by line before it could claim a
// slot (ncu r1_k: 7.6 GB of DRAM traffic for 2.8 GB of algorithmic bytes, the look + claim chain of every row waiting on
// DRAM). Here the work items of ONE ordered sequence are handed out by a global counter:
//     C(0), C(1) I(0), C(2) I(1), ... C(P-1) I(P-2), I(P-1)        C(p) = clear partition p's table (256 KB per item)
//                                                                   I(p) = insert 1024 rows of partition p
// so a partition's table is zeroed (full-line writes: no DRAM read) just before its rows arrive and is still in L2 when they
// do. An I(p) item waits until every C(p) item has FINISHED (done[p] == its item count): all of them were claimed earlier in
// the sequence by CTAs that are running, so the wait is bounded; a watchdog gives up (error flag) instead of hanging.
// ------------------------------------------------------------------------------------------------
constexpr uint32_t FB_CLEAR_BYTES = 256u << 10;
constexpr uint32_t FB_TILE_ROWS = 1024;
struct FusedBuildArgs {
    BuildArgs           b;            // keys / rowid = the partitioned build stream, n = its rows
    const uint32_t     *part_off;     // [P + 1] first row of every partition in the stream
    const uint32_t     *seq_off;      // [2P + 1] first item of every group of the sequence: group 2p = C(p), group 2p + 1 = I(p) (see fused_group)
    uint32_t            n_parts;
    uint32_t            n_items;
    unsigned int       *done;         // [P] finished clear items per partition
    unsigned int       *error;        // watchdog
};
// group g of the sequence: g = 0 -> C(0); then pairs (C(p + 1), I(p)) for p = 0 .. P-2; last group I(P-1)
__device__ __forceinline__ void fused_group(uint32_t g, uint32_t P, bool &is_clear, uint32_t &p) {
    if (g == 0) { is_clear = true; p = 0; return; }
    const uint32_t k = g - 1;                         // k even: C(k/2 + 1), k odd: I(k/2)   -- except past the last clear
    if (k < 2 * (P - 1)) { is_clear = (k & 1u) == 0u; p = is_clear ? k / 2 + 1 : k / 2; return; }
    is_clear = false; p = P - 1;
}

template <int KW>
__global__ void __launch_bounds__(TILE_THREADS) build_insert_fused_kernel(FusedBuildArgs f) {
    using KeyT = typename std::conditional<KW == 8, unsigned long long, uint32_t>::type;
    const BuildArgs &a = f.b;
    const int t = threadIdx.x;
    const uint32_t P = f.n_parts, n_groups = 2 * P;
    uint32_t distinct = 0, dups = 0, multi = 0;
    __shared__ unsigned int s_item;
    for (;;) {
        __syncthreads();
        if (t == 0) s_item = atomicAdd(&a.ctr->next_tile, 1u);
        __syncthreads();
        const uint32_t item = s_item;
        if (item >= f.n_items) break;
        // group of this item: the last g with seq_off[g] <= item (2P groups: binary search)
        uint32_t lo = 0, hi = n_groups;
        while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (__ldg(f.seq_off + mid) <= item) lo = mid; else hi = mid; }
        bool is_clear; uint32_t p;
        fused_group(lo, P, is_clear, p);
        const uint32_t local = item - __ldg(f.seq_off + lo);
        const uint2 pe = __ldg(reinterpret_cast<const uint2 *>(a.tv.parts + p));          // {base bucket, buckets}
        if (is_clear) {
            const unsigned long long bytes = (unsigned long long)pe.y * 32ull;
            const unsigned long long b0 = (unsigned long long)local * FB_CLEAR_BYTES, b1 = min(b0 + FB_CLEAR_BYTES, bytes);
            char *tbl = reinterpret_cast<char *>(a.tv.slots) + (unsigned long long)pe.x * 32ull;
            for (unsigned long long o = b0 + (unsigned long long)t * 16ull; o < b1; o += (unsigned long long)TILE_THREADS * 16ull)
                *reinterpret_cast<uint4 *>(tbl + o) = make_uint4(0u, 0u, 0u, 0u);
            __syncthreads();
            if (t == 0) { __threadfence(); atomicAdd(f.done + p, 1u); }
            continue;
        }
        // ---- I(p): wait for the partition's table to be zeroed
        if (t == 0) {
            const uint32_t need = (uint32_t)(((unsigned long long)pe.y * 32ull + FB_CLEAR_BYTES - 1) / FB_CLEAR_BYTES);
            unsigned int spins = 0;
            while (*reinterpret_cast<volatile unsigned int *>(f.done + p) < need) {
                __nanosleep(200);
                if (++spins > (1u << 23)) { atomicExch(f.error, 1u); break; }           // ~2 s: never in a correct run
            }
            __threadfence();
        }
        __syncthreads();
        const uint32_t r0 = __ldg(f.part_off + p) + local * FB_TILE_ROWS, r1 = min(r0 + FB_TILE_ROWS, __ldg(f.part_off + p + 1));
        const KeyT *keys = reinterpret_cast<const KeyT *>(a.keys);
#pragma unroll
        for (int q = 0; q < (int)(FB_TILE_ROWS / TILE_THREADS); q++) {
            const uint32_t row = r0 + (uint32_t)q * TILE_THREADS + t;
            if (row >= r1) continue;
            const uint64_t key = (uint64_t)__ldcs(keys + row);
            const uint32_t r = __ldcs(a.rowid + row);
            const uint32_t d0 = distinct;
            Slot *s = find_or_claim<true, true>(a.tv, a.hk, key, distinct, r);
            if (a.slot_of) a.slot_of[row] = key ? (uint32_t)(s - a.tv.slots) : 0xffffffffu;
            if (distinct != d0) continue;
            uint32_t pos;
            const unsigned am = __activemask();
            if (__any_sync(am, distinct == d0)) {
                const unsigned peers = __match_any_sync(am, (unsigned long long)s);
                const int leader = __ffs(peers) - 1;
                uint32_t base = 0;
                if ((t & 31) == leader) base = atomicAdd(&s->count, (uint32_t)__popc(peers)) & COUNT_MASK;
                base = __shfl_sync(peers, base, leader);
                pos = base + __popc(peers & ((1u << (t & 31)) - 1u));
            } else {
                pos = atomicAdd(&s->count, 1u) & COUNT_MASK;
            }
            if (pos == 0) {
                s->row = r;
                if (key == 0) distinct++;
            } else {
                dups++;
                if (pos == 1) multi++;
            }
        }
    }
    const unsigned long long d = warp_sum_u32(distinct), u = warp_sum_u32(dups), m = warp_sum_u32(multi);
    if ((t & 31) == 0) {
        if (d) atomicAdd(&a.ctr->distinct, d);
        if (u) atomicAdd(&a.ctr->dup_rows, u);
        if (m) atomicAdd(&a.ctr->multi_slots, m);
    }
}

// Duplicate keys: turn every slot with count >= 2 into a run header. The run takes count + 1 entries of
// rowids[]: [cursor][row 0][row 1]...; slot.row points at "row 0". One warp-aggregated claim per warp.
__global__ void __launch_bounds__(256) csr_scan_kernel(TableView tv, uint64_t n_slots, Counters *ctr) {
    const int lane = threadIdx.x & 31;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const uint64_t start = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    // n_slots + 1: the extra index is the zero-key side slot
    for (uint64_t i0 = start - lane; i0 < n_slots + 1; i0 += stride) {
        const uint64_t i = i0 + lane;
        Slot *s = nullptr;
        uint32_t need = 0;
        if (i < n_slots) s = tv.slots + i; else if (i == n_slots) s = tv.zero_slot;
        Slot v{0, 0, 0};
        if (s) { v = load_slot(s); if (v.count >= 2) need = v.count + 1; }
        // warp exclusive scan of `need`
        uint32_t incl = need;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
        const uint32_t tot = __shfl_sync(0xffffffffu, incl, 31);
        unsigned long long base = 0;
        if (tot) {
            if (lane == 0) base = atomicAdd(&ctr->csr_cursor, (unsigned long long)tot);
            base = __shfl_sync(0xffffffffu, base, 0);
        }
        if (need) {
            const uint32_t off = (uint32_t)(base + incl - need);
            tv.rowids[off] = 1u;            // fill cursor: position 0 is taken by the first row
            tv.rowids[off + 1] = v.row;
            s->row = off + 1;
        }
    }
}

template <int KW, bool RADIX>
__global__ void __launch_bounds__(TILE_THREADS) csr_fill_kernel(BuildArgs a) {
    const int t = threadIdx.x;
    const int64_t n_tiles = (a.n + a.mis + TILE_ROWS - 1) / TILE_ROWS;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        uint64_t key[TILE_ITEMS];
        int64_t row[TILE_ITEMS];
        load_tile_keys<KW>(a.keys, a.mis, a.n, tile, t, key, row);
#pragma unroll
        for (int q = 0; q < TILE_ITEMS; q++) {
            if (row[q] < 0 || row[q] >= a.n) continue;
            if (!bit_valid(a.validity, row[q])) continue;
            const uint32_t r = a.rowid ? __ldcs(a.rowid + row[q]) : (uint32_t)row[q];
            Slot *s;
            if (a.slot_of) {                                                 // the insert pass remembered where the row went
                uint32_t si = __ldcs(a.slot_of + row[q]);
                if (a.dup_bit && si != 0xffffffffu) {
                    if (!(si & SLOT_OF_DUP)) continue;                       // the row claimed its slot itself: placed by csr_scan if the slot became a run
                    si &= ~SLOT_OF_DUP;
                }
                s = si == 0xffffffffu ? a.tv.zero_slot : a.tv.slots + si;
            } else {
                uint32_t dummy = 0;
                s = find_or_claim<RADIX>(a.tv, a.hk, key[q], dummy);         // key is present: never claims
            }
            const Slot v = load_slot(s);
            if (v.count < 2) continue;
            if (__ldcg(a.tv.rowids + v.row) == r) continue;                  // the run's first row is already placed
            const unsigned am = __activemask();                              // lanes of the same run: one cursor atomic
            const unsigned peers = __match_any_sync(am, v.row);
            const int leader = __ffs(peers) - 1;
            uint32_t base = 0;
            if ((t & 31) == leader) base = atomicAdd(a.tv.rowids + v.row - 1, (uint32_t)__popc(peers));
            base = __shfl_sync(peers, base, leader);
            a.tv.rowids[v.row + base + __popc(peers & ((1u << (t & 31)) - 1u))] = r;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// a4 + a5 + a7: probe
// ------------------------------------------------------------------------------------------------
enum ProbeMode : int {
    PM_PAIRS = 0,   // Inner / Left: (build, probe) for every match
    PM_OUTER = 1,   // Right / Full: + (NULL, probe) for probe rows without a match
    PM_SEMI  = 2,   // RightSemi: probe rows with a match, once
    PM_ANTI  = 3,   // RightAnti: probe rows without a match
    PM_MARK  = 4,   // LeftSemi / LeftAnti: only mark visited build rows
};

// A duplicate run too long for one warp (a heavy-hitter key: C3's Zipf build side has runs of millions of rows). The probe
// kernel reserves the run's output range as usual but leaves the copy to expand_runs_kernel, where the whole grid shares
// it: with the copy inside the probe kernel, ONE warp walked a 2.7 M-row run for every probe row that hit the key and the
// launch took 135 ms for 1 B probe rows; deferred, it is bounded by the output bandwidth.
struct BigRun {
    unsigned long long pos;       // first output row of the run's (build, probe) pairs
    uint32_t           run;       // start of the run in rowids[]
    uint32_t           cnt;
    uint32_t           prow;      // the probe row
    uint32_t           _pad;
};
constexpr uint32_t BIG_RUN_ROWS = 4096;

struct ProbeArgs {
    const void     *keys;
    const uint8_t  *validity;
    int64_t         bit_offset;
    const uint32_t *rowid;        // explicit probe row ids (radix layout) or null = identity
    int64_t         n;            // < 2^32 - 16
    int32_t         mis;
    int32_t         _pad;
    TableView       tv;
    HashKeys        hk;
    uint64_t       *out_build;    // may be null (semi / anti / mark)
    uint32_t       *out_probe;
    uint64_t        out_cap;
    uint32_t       *visited;      // bitmap over build rows (MARK)
    struct BigRun  *big_runs;     // list for duplicate runs too long for one warp (or null)
    uint32_t        big_cap;
    uint32_t        _pad2;
    Counters       *ctr;
};

#ifndef B200J_PROBE_THREADS
#define B200J_PROBE_THREADS 128
#endif
#ifndef B200J_PROBE_ITEMS
#define B200J_PROBE_ITEMS 4
#endif
constexpr int PROBE_THREADS = B200J_PROBE_THREADS;
constexpr int PROBE_ITEMS   = B200J_PROBE_ITEMS;
constexpr int PROBE_TILE    = PROBE_THREADS * PROBE_ITEMS;

__device__ __forceinline__ void mark_visited(uint32_t *visited, uint32_t row) {
    uint32_t *w = visited + (row >> 5);
    const uint32_t bit = 1u << (row & 31);
    if (!(__ldcg(w) & bit)) atomicOr(w, bit);
}

// One long duplicate run (>= 32 rows), called by the whole warp with warp-uniform arguments: deferred to the grid-wide
// expansion when it is a heavy hitter, else copied here, coalesced.
template <bool EMIT, bool MARK>
__device__ __forceinline__ void copy_long_run(const ProbeArgs &a, unsigned long long pos, uint32_t cnt, uint32_t rb, uint32_t pr, int lane) {
    if (cnt >= BIG_RUN_ROWS && a.big_runs != nullptr) {
        uint32_t slot = 0;
        if (lane == 0) slot = atomicAdd(&a.ctr->big_count, 1u);
        slot = __shfl_sync(0xffffffffu, slot, 0);
        if (slot < a.big_cap) {
            if (lane == 0) a.big_runs[slot] = BigRun{pos, rb, cnt, pr, 0u};
            return;
        }
    }
    for (uint32_t k = lane; k < cnt; k += 32) {
        const uint32_t br = __ldcg(a.tv.rowids + rb + k);
        if (EMIT && pos + k < a.out_cap) {
            a.out_build[pos + k] = (uint64_t)br;
            a.out_probe[pos + k] = pr;
        }
        if (MARK) mark_visited(a.visited, br);
    }
}

// The deferred runs, shared out over the grid in chunks of EXPAND_CHUNK rows (global chunk g goes to CTA g % gridDim.x).
constexpr uint32_t EXPAND_CHUNK = 2048;
template <bool EMIT, bool MARK>
__global__ void __launch_bounds__(256) expand_runs_kernel(const BigRun *__restrict__ runs, uint32_t n_runs, const uint32_t *__restrict__ rowids,
                                                          uint64_t *__restrict__ out_build, uint32_t *__restrict__ out_probe, uint64_t out_cap,
                                                          uint32_t *visited) {
    unsigned long long before = 0;                       // chunks of the runs in front of this one
    for (uint32_t d = 0; d < n_runs; d++) {
        const BigRun r = runs[d];
        const uint32_t chunks = (r.cnt + EXPAND_CHUNK - 1) / EXPAND_CHUNK;
        const uint32_t first = (uint32_t)((blockIdx.x + gridDim.x - (uint32_t)(before % gridDim.x)) % gridDim.x);
        for (uint32_t c = first; c < chunks; c += gridDim.x) {
            const uint32_t k1 = min(r.cnt, (c + 1u) * EXPAND_CHUNK);
            for (uint32_t k = c * EXPAND_CHUNK + threadIdx.x; k < k1; k += blockDim.x) {
                const uint32_t br = __ldcs(rowids + r.run + k);
                if (EMIT && r.pos + k < out_cap) {
                    __stcs(reinterpret_cast<unsigned long long *>(out_build) + r.pos + k, (unsigned long long)br);
                    __stcs(out_probe + r.pos + k, r.prow);
                }
                if (MARK) mark_visited(visited, br);
            }
        }
        before += chunks;
    }
}

// Keys of one tile. GUARD = false: the tile lies fully inside the batch and there is no validity bitmap, so
// there is not a single range or NULL check on the hot path. `in` / `look`: bit q set = item q is a row of the
// batch / is a row with a non-NULL key.
template <int KW, bool GUARD>
__device__ __forceinline__ void probe_load_tile(const ProbeArgs &a, uint32_t tile, int t, uint64_t (&key)[PROBE_ITEMS],
                                                uint32_t (&row)[PROBE_ITEMS], uint32_t &in, uint32_t &look) {
    constexpr int PER_VEC = 16 / KW;
    constexpr int NVEC = PROBE_ITEMS / PER_VEC;
    const uint32_t vbase = tile * (uint32_t)PROBE_TILE;
    const uint32_t mis = (uint32_t)a.mis, vend = (uint32_t)a.n + mis;
    in = look = (PROBE_ITEMS == 32) ? 0xffffffffu : ((1u << PROBE_ITEMS) - 1u);
    if (KW == 8) {
        const unsigned long long *k64 = reinterpret_cast<const unsigned long long *>(a.keys) - mis;
#pragma unroll
        for (int v = 0; v < NVEC; v++) {
            const uint32_t vi = vbase + ((uint32_t)v * PROBE_THREADS + t) * 2u;
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
            const uint32_t vi = vbase + ((uint32_t)v * PROBE_THREADS + t) * 4u;
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
    look = in;
    if (GUARD && a.validity != nullptr) {
#pragma unroll
        for (int q = 0; q < PROBE_ITEMS; q++)
            if (((in >> q) & 1u) && !bit_valid(a.validity, a.bit_offset + (int64_t)row[q])) look &= ~(1u << q);
    }
}

#ifndef B200J_PROBE_SUB
#define B200J_PROBE_SUB 4
#endif
constexpr int PROBE_SUB   = B200J_PROBE_SUB;                 // tiles per output flush
constexpr int PROBE_STAGE = PROBE_TILE * PROBE_SUB;          // rows staged in shared memory per global atomic

struct ProbeStage {                                          // block-local staging of single-row results
    uint32_t brow[PROBE_STAGE];                              // build row (0xffffffff = NULL)
    uint32_t prow[PROBE_STAGE];
    uint32_t count;
    uint32_t next;                                           // next super-tile of this CTA
    unsigned long long gbase;
};

// One tile of PROBE_TILE rows. No block barrier in here: warps run free inside a super-tile; rows that produce
// exactly one output go to the shared-memory stage (one warp-aggregated shared atomic per warp), duplicate-key
// runs (count >= 2) reserve their own global range with one warp-aggregated global atomic.
template <int KW, int MODE, bool MARK, bool RADIX, bool GUARD>
__device__ __forceinline__ void probe_tile(const ProbeArgs &a, uint32_t tile, int t, ProbeStage &st, uint32_t &null_build_out) {
    constexpr uint32_t FULL = 0xffffffffu;
    const int lane = t & 31;
    const uint32_t lt_mask = (1u << lane) - 1u;

    uint64_t key[PROBE_ITEMS];
    uint32_t row[PROBE_ITEMS];
    uint32_t in, look;
    probe_load_tile<KW, GUARD>(a, tile, t, key, row, in, look);

    // ---- round 0: hash + bucket load of every item before any is consumed (PROBE_ITEMS random 32 B loads in flight).
    // The key value 0 lives in the side bucket (whose key word is 0 as well), so the same compare serves both.
    uint32_t idx[PROBE_ITEMS], hp[PROBE_ITEMS];               // idx = bucket inside the partition
    uint32_t srow[PROBE_ITEMS], scnt[PROBE_ITEMS];          // the matching slot (scnt == 0: no match)
    uint32_t pending = GUARD ? look : in;
    {
        unsigned long long k0[PROBE_ITEMS], v0[PROBE_ITEMS], k1[PROBE_ITEMS], v1[PROBE_ITEMS];
#pragma unroll
        for (int q = 0; q < PROBE_ITEMS; q++) {
            const uint64_t h = hash_one(a.hk, key[q]);
            uint32_t base, nb;
            hp[q] = (uint32_t)(h >> 32);
            locate<RADIX>(a.tv, hp[q], base, nb);
            idx[q] = home_bucket(h, nb);
            const Slot *p = key[q] ? bucket_ptr(a.tv, base, idx[q]) : a.tv.zero_slot;
            k0[q] = v0[q] = k1[q] = v1[q] = 0ULL;
            if (!GUARD || ((pending >> q) & 1u)) load_bucket(p, k0[q], v0[q], k1[q], v1[q], (!RADIX && a.tv.l1 != 0u));
        }
#pragma unroll
        for (int q = 0; q < PROBE_ITEMS; q++) {
            const bool hit0 = k0[q] == key[q], hit1 = k1[q] == key[q];
            const bool done = hit0 | hit1 | (k0[q] == 0ULL) | (k1[q] == 0ULL) | !bucket_overflowed(v0[q]);   // home bucket: see SLOT_OVF
            const unsigned long long v = hit0 ? v0[q] : ((hit1 && k0[q] != 0ULL) ? v1[q] : 0ULL);
            srow[q] = (uint32_t)v;
            scnt[q] = (uint32_t)(v >> 32) & COUNT_MASK;
            if (done) pending &= ~(1u << q);
        }
    }
    // ---- later rounds (about one lookup in ten at load factor 0.5): each lane walks its lowest unresolved item
    while (pending) {
        const int q = __ffs(pending) - 1;
        uint64_t k = key[0];
        uint32_t i = idx[0], p = hp[0];
#pragma unroll
        for (int j = 1; j < PROBE_ITEMS; j++)
            if (q == j) { k = key[j]; i = idx[j]; p = hp[j]; }
        uint32_t base, nb;
        locate<RADIX>(a.tv, p, base, nb);
        for (;;) {
            i = next_bucket(i, nb);
            unsigned long long k0, v0, k1, v1;
            load_bucket(bucket_ptr(a.tv, base, i), k0, v0, k1, v1, (!RADIX && a.tv.l1 != 0u));
            const bool hit0 = k0 == k, hit1 = k1 == k;
            if (hit0 | hit1 | (k0 == 0ULL) | (k1 == 0ULL)) {
                const unsigned long long v = hit0 ? v0 : ((hit1 && k0 != 0ULL) ? v1 : 0ULL);
#pragma unroll
                for (int j = 0; j < PROBE_ITEMS; j++)
                    if (q == j) { srow[j] = (uint32_t)v; scnt[j] = (uint32_t)(v >> 32) & COUNT_MASK; }
                break;
            }
        }
        pending &= pending - 1u;
    }

    // ---- classify: `one` = items producing exactly one staged row, `multi` = duplicate-key runs
    uint32_t one = 0, multi = 0, found = 0;
#pragma unroll
    for (int q = 0; q < PROBE_ITEMS; q++) {
        const bool f = (!GUARD || ((look >> q) & 1u)) && scnt[q] != 0u;
        const bool inq = !GUARD || ((in >> q) & 1u);
        if (f) found |= 1u << q;
        if (MODE == PM_PAIRS) { if (f && scnt[q] == 1u) one |= 1u << q; if (f && scnt[q] > 1u) multi |= 1u << q; }
        else if (MODE == PM_OUTER) { if ((f && scnt[q] == 1u) || (inq && !f)) one |= 1u << q; if (f && scnt[q] > 1u) multi |= 1u << q; }
        else if (MODE == PM_SEMI) { if (f) one |= 1u << q; }
        else if (MODE == PM_ANTI) { if (inq && !f) one |= 1u << q; }
        if (MARK && MODE != PM_PAIRS && MODE != PM_OUTER && f && scnt[q] > 1u) multi |= 1u << q;
    }

    // ---- single rows -> shared-memory stage: ballot positions, one shared atomic per warp
    if (MODE != PM_MARK) {
        uint32_t off[PROBE_ITEMS];
        uint32_t wtot = 0;
#pragma unroll
        for (int q = 0; q < PROBE_ITEMS; q++) {
            const uint32_t b = __ballot_sync(FULL, (one >> q) & 1u);
            off[q] = wtot + __popc(b & lt_mask);
            wtot += __popc(b);
        }
        uint32_t wbase = 0;
        if (lane == 0 && wtot) wbase = atomicAdd(&st.count, wtot);
        wbase = __shfl_sync(FULL, wbase, 0);
#pragma unroll
        for (int q = 0; q < PROBE_ITEMS; q++) {
            if (!((one >> q) & 1u)) continue;
            const bool f = (found >> q) & 1u;
            const bool null_build = !f || MODE == PM_SEMI || MODE == PM_ANTI;
            st.brow[wbase + off[q]] = null_build ? 0xffffffffu : srow[q];
            st.prow[wbase + off[q]] = a.rowid ? __ldcs(a.rowid + row[q]) : row[q];
            if (null_build) null_build_out++;
            else if (MARK) mark_visited(a.visited, srow[q]);
        }
    } else {
#pragma unroll
        for (int q = 0; q < PROBE_ITEMS; q++)
            if (((found >> q) & 1u) && scnt[q] == 1u) mark_visited(a.visited, srow[q]);
    }

    // ---- duplicate-key runs (rare unless the build side is skewed): own global reservation per warp
    if (__any_sync(FULL, multi != 0u)) {
        const bool emit = (MODE == PM_PAIRS || MODE == PM_OUTER);
        uint32_t off[PROBE_ITEMS];
        uint32_t wtot = 0;
        if (emit) {
#pragma unroll
            for (int q = 0; q < PROBE_ITEMS; q++) {
                const uint32_t mq = ((multi >> q) & 1u) ? scnt[q] : 0u;
                uint32_t incl = mq;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += y; }
                off[q] = wtot + incl - mq;
                wtot += __shfl_sync(FULL, incl, 31);
            }
        }
        unsigned long long gbase = 0;
        if (emit) {
            if (lane == 0) gbase = atomicAdd(&a.ctr->out_count, (unsigned long long)wtot);
            gbase = __shfl_sync(FULL, gbase, 0);
        }
#pragma unroll
        for (int q = 0; q < PROBE_ITEMS; q++) {
            const bool mine = (multi >> q) & 1u;
            const uint32_t prow = (mine && a.rowid) ? __ldcs(a.rowid + row[q]) : row[q];
            // short runs: the owning lane copies; long runs: the whole warp copies one run at a time (coalesced)
            if (mine && scnt[q] < 32u) {
                const unsigned long long pos = gbase + off[q];
                for (uint32_t k = 0; k < scnt[q]; k++) {
                    const uint32_t br = __ldcg(a.tv.rowids + srow[q] + k);
                    if (emit && pos + k < a.out_cap) {
                        a.out_build[pos + k] = (uint64_t)br;
                        a.out_probe[pos + k] = prow;
                    }
                    if (MARK) mark_visited(a.visited, br);
                }
            }
            uint32_t pend = __ballot_sync(FULL, mine && scnt[q] >= 32u);
            while (pend) {
                const int src = __ffs(pend) - 1;
                pend &= pend - 1;
                const unsigned long long pos = __shfl_sync(FULL, gbase + (emit ? off[q] : 0u), src);
                const uint32_t cnt = __shfl_sync(FULL, scnt[q], src);
                const uint32_t rb = __shfl_sync(FULL, srow[q], src);
                const uint32_t pr = __shfl_sync(FULL, prow, src);
                copy_long_run<(MODE == PM_PAIRS || MODE == PM_OUTER), MARK>(a, pos, cnt, rb, pr, lane);
            }
        }
    }
}

template <int KW, int MODE, bool MARK, bool RADIX>
__global__ void __launch_bounds__(PROBE_THREADS) probe_kernel(ProbeArgs a) {
    __shared__ ProbeStage st;
    const int t = threadIdx.x;
    const uint32_t mis = (uint32_t)a.mis;
    const uint32_t vend = (uint32_t)a.n + mis;
    const uint32_t n_tiles = (vend + PROBE_TILE - 1) / PROBE_TILE;
    const uint32_t n_super = (n_tiles + PROBE_SUB - 1) / PROBE_SUB;
    const bool no_nulls = a.validity == nullptr;
    uint32_t null_build_out = 0;
    // Super-tiles are handed out by a global counter, not by a static grid stride: the running CTAs then always
    // work on the most recent ~1200 super-tiles, i.e. on one or two radix partitions, whose tables stay in L2.
    // (With a static stride the CTAs drift apart by dozens of partitions and every lookup goes to HBM.)
    if (t == 0) { st.count = 0; st.next = atomicAdd(&a.ctr->next_tile, 1u); }
    __syncthreads();
    for (;;) {
        const uint32_t sup = st.next;
        if (sup >= n_super) break;
#pragma unroll 1
        for (int sub = 0; sub < PROBE_SUB; sub++) {
            const uint32_t tile = sup * PROBE_SUB + sub;
            if (tile >= n_tiles) break;
            const bool inner = no_nulls && (tile > 0 || mis == 0) && (tile + 1) * (uint32_t)PROBE_TILE <= vend;
            if (inner) probe_tile<KW, MODE, MARK, RADIX, false>(a, tile, t, st, null_build_out);
            else probe_tile<KW, MODE, MARK, RADIX, true>(a, tile, t, st, null_build_out);
        }
        // ---- flush: one global atomic per super-tile, then fully coalesced streaming 8 B + 4 B stores
        __syncthreads();
        const uint32_t cnt = st.count;
        if (t == 0) st.gbase = cnt ? atomicAdd(&a.ctr->out_count, (unsigned long long)cnt) : 0ULL;
        __syncthreads();
        const unsigned long long gbase = st.gbase;
        if (gbase + cnt <= a.out_cap) {
            if (a.out_build)
                for (uint32_t i = t; i < cnt; i += PROBE_THREADS) {
                    const uint32_t br = st.brow[i];
                    __stcs(reinterpret_cast<unsigned long long *>(a.out_build) + gbase + i, br == 0xffffffffu ? ~0ULL : (unsigned long long)br);
                }
            for (uint32_t i = t; i < cnt; i += PROBE_THREADS) __stcs(a.out_probe + gbase + i, st.prow[i]);
        } else {
            for (uint32_t i = t; i < cnt; i += PROBE_THREADS) {
                if (gbase + i >= a.out_cap) break;
                const uint32_t br = st.brow[i];
                if (a.out_build) a.out_build[gbase + i] = br == 0xffffffffu ? ~0ULL : (uint64_t)br;
                a.out_probe[gbase + i] = st.prow[i];
            }
        }
        __syncthreads();
        if (t == 0) { st.count = 0; st.next = atomicAdd(&a.ctr->next_tile, 1u); }
        __syncthreads();
    }
    if (MODE == PM_OUTER || MODE == PM_SEMI || MODE == PM_ANTI) {
        const unsigned long long z = warp_sum_u32(null_build_out);
        if ((t & 31) == 0 && z) atomicAdd(&a.ctr->build_null_out, z);
    }
}

// ------------------------------------------------------------------------------------------------
// a7 helpers: bitmap -> indices, pairs -> bitmaps
// ------------------------------------------------------------------------------------------------
// Writes the indices i < n whose bit equals want_set, as u64 (out64) or u32 (out32).
__global__ void __launch_bounds__(256) bitmap_select_kernel(const uint32_t *__restrict__ bm, int64_t n, int want_set,
                                                            uint64_t *out64, uint32_t *out32, uint64_t out_cap,
                                                            unsigned long long *counter) {
    __shared__ uint32_t s_warp[8];
    __shared__ unsigned long long s_base;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int64_t n_words = (n + 31) / 32;
    const int64_t n_iter = (n_words + blockDim.x - 1) / blockDim.x;
    for (int64_t it = blockIdx.x; it < n_iter; it += gridDim.x) {
        const int64_t w = it * blockDim.x + t;
        uint32_t bits = 0;
        if (w < n_words) {
            bits = bm[w];
            if (!want_set) bits = ~bits;
            const int64_t rem = n - w * 32;
            if (rem < 32) bits &= (1u << rem) - 1u;
        }
        const uint32_t c = __popc(bits);
        uint32_t incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        if (t == 0) {
            unsigned long long tot = 0;
            for (int i = 0; i < 8; i++) tot += s_warp[i];
            s_base = tot ? atomicAdd(counter, tot) : 0ULL;
        }
        __syncthreads();
        unsigned long long pos = s_base + incl - c;
        for (int i = 0; i < warp; i++) pos += s_warp[i];
        __syncthreads();
        while (bits) {
            const int b = __ffs(bits) - 1;
            bits &= bits - 1;
            if (pos < out_cap) {
                if (out64) out64[pos] = (uint64_t)(w * 32 + b);
                if (out32) out32[pos] = (uint32_t)(w * 32 + b);
            }
            pos++;
        }
    }
}

// apply_pairs: mark `seen` probe rows and `visited` build rows of caller-filtered pairs.
// Indices come from the caller: a probe index >= n_probe_rows or a build index >= n_build (other than the NULL marker) is
// counted in *n_bad and marks nothing.
__global__ void __launch_bounds__(256) mark_pairs_kernel(const uint64_t *__restrict__ build_idx, const uint32_t *__restrict__ probe_idx,
                                                         int64_t n, uint32_t *seen, uint32_t *visited, uint64_t n_probe_rows, uint64_t n_build,
                                                         unsigned long long *n_bad) {
    uint32_t bad = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const uint64_t b = build_idx[i];
        const uint32_t p = probe_idx[i];
        if ((uint64_t)p >= n_probe_rows || (b != ~0ULL && b >= n_build)) { bad++; continue; }
        if (seen) mark_visited(seen, p);
        if (visited && b != ~0ULL) mark_visited(visited, (uint32_t)b);
    }
    const unsigned long long z = warp_sum_u32(bad);
    if ((threadIdx.x & 31) == 0 && z) atomicAdd(n_bad, z);
}

// Fills (NULL, probe) / (build, NULL) sides of index-only results.
__global__ void __launch_bounds__(256) fill_u64_kernel(uint64_t *p, int64_t n, uint64_t v) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}
__global__ void __launch_bounds__(256) fill_u32_kernel(uint32_t *p, int64_t n, uint32_t v) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}

// B200J_OUT_BUILD_U32: build indices narrowed to 32 bits (the NULL marker ~0 stays all ones)
__global__ void __launch_bounds__(256) narrow_u64_u32_kernel(const uint64_t *__restrict__ in, uint32_t *__restrict__ out, int64_t n) {
    const int64_t n2 = n >> 1;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += (int64_t)gridDim.x * blockDim.x) {
        const ulonglong2 v = __ldcs(reinterpret_cast<const ulonglong2 *>(in) + i);
        __stcs(reinterpret_cast<uint2 *>(out) + i, make_uint2((uint32_t)v.x, (uint32_t)v.y));
    }
    if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) out[n - 1] = (uint32_t)in[n - 1];
}

// B200J_OUT_PROBE_ORDER: pairs (one per probe row, unique build keys) -> ONE array in the order of the probe batch,
// out[probe_idx - base] = build_idx. The pairs of the dense probes come in the order of the partitioned stream, i.e. the probe
// indices of 32 K consecutive pairs stay inside one 32 K-row window of `out`: the scattered 4 / 8 B writes merge in L2.
template <typename OutT>
__global__ void __launch_bounds__(256) reorder_by_probe_kernel(const uint64_t *__restrict__ bidx, const uint32_t *__restrict__ pidx, int64_t n,
                                                               uint32_t base, OutT *__restrict__ out) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const unsigned long long b = __ldcs(reinterpret_cast<const unsigned long long *>(bidx) + i);
        const uint32_t p = __ldcs(pidx + i) - base;
        if ((int64_t)p < n) out[p] = (OutT)b;               // ~0ULL narrows to the 32-bit NULL index
    }
}

// 32-bit key pieces of the multi-GPU exchange -> 8-byte keys (the concatenating fall-back of b200j_probe_pieces)
__global__ void __launch_bounds__(256) widen_u32_u64_kernel(const uint32_t *__restrict__ in, unsigned long long *__restrict__ out, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = (unsigned long long)in[i];
}

// Probe row ids of one chunk of a host batch -> row ids of the batch (direct layout; the radix layout's partition pass
// adds the base while it writes the row-id array). NULL markers stay.
__global__ void __launch_bounds__(256) add_base_u32_kernel(uint32_t *p, int64_t n, uint32_t base) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t v = p[i];
        if (v != 0xffffffffu) p[i] = v + base;
    }
}

// Validity concat for the build store: dst bits [dst_off, dst_off + n) <- src bits [src_off, src_off + n)
// (src == null: all ones). One thread per destination word; words are touched by one thread only.
__global__ void __launch_bounds__(256) copy_bits_kernel(uint32_t *dst, int64_t dst_off, const uint8_t *src, int64_t src_off, int64_t n) {
    const int64_t w0 = dst_off >> 5, w1 = (dst_off + n + 31) >> 5;
    for (int64_t w = w0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < w1; w += (int64_t)gridDim.x * blockDim.x) {
        uint32_t bits = 0, msk = 0;
        for (int b = 0; b < 32; b++) {
            const int64_t g = w * 32 + b - dst_off;
            if (g < 0 || g >= n) continue;
            msk |= 1u << b;
            const int64_t sidx = src_off + g;
            if (src == nullptr || ((src[sidx >> 3] >> (sidx & 7)) & 1)) bits |= 1u << b;
        }
        dst[w] = (dst[w] & ~msk) | bits;
    }
}

}  // namespace b200j
