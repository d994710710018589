/*
 * oracle/join_oracle.c — CPU restatement of the reference's hash-join hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under datafusion_parallelism_b200/ may link, import or
 * call this file; it is the checker for tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs.  It is deliberately scalar and simple.
 *
 * Reference: jamesfer/datafusion-parallelism @ 3e184de (Rust, not buildable here: no cargo,
 * nightly-only, version10 is aarch64-only).  Every function cites the reference file:line it
 * restates.  Third-party arithmetic that is NOT under /root/reference and is restated from its
 * published algorithm:
 *   - ahash 0.8.11 (Cargo.lock:21-32), fallback hasher + `specialize` integer path,
 *     RandomState::with_seed(0) over the fixed PI/PI2 seeds (neither rng feature);
 *   - datafusion-common 40.0.0 hash_utils::create_hashes / combine_hashes.
 *
 * PARITY STATUS
 *   - join RESULTS: pinned against the literal expected batches of the reference's own tests
 *     (src/lib.rs:97-131,185-192,244-245,299-306,366-371,431-436,490-497,568-573,626-633,686-725),
 *     the chain-order KATs of the map unit tests, the bitset KATs and the generator KAT —
 *     see tests/test_oracle_golden.py and tests/golden/.
 *   - HASH VALUES: "parity unpinned".  No reference test pins a hash value and the value is not
 *     even well defined across builds of the reference (ahash compile-time-rng).  The hash is a
 *     parameterised exact function (keys[4]); the contract is GPU hash == this oracle's hash
 *     bit for bit.  Join results never depend on it (equal_rows_arr compares key values).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>

#define ORC_NULL64 UINT64_MAX
#define ORC_NULL32 UINT32_MAX

/* DataFusion JoinType discriminants (datafusion-common 40 join_type.rs order). */
enum { ORC_INNER = 0, ORC_LEFT = 1, ORC_RIGHT = 2, ORC_FULL = 3,
       ORC_LEFT_SEMI = 4, ORC_RIGHT_SEMI = 5, ORC_LEFT_ANTI = 6, ORC_RIGHT_ANTI = 7 };

/* One Arrow key column: values buffer, optional LSB-first validity bitmap; Utf8 adds 32-bit offsets. */
typedef struct {
    const void    *values;     /* fixed-width values, or the Utf8 data bytes */
    const uint8_t *validity;   /* NULL => all valid */
    int32_t        width;      /* 4 (Int32/UInt32/Date32), 8 (Int64/UInt64/…) or 0 = Utf8 */
    int32_t        _pad;
    const int32_t *offsets;    /* Utf8: n_rows + 1 offsets into values */
} orc_col;

/* ------------------------------------------------------------------------------------------ */
/* ahash 0.8.11 fallback arithmetic                                                           */
/* ------------------------------------------------------------------------------------------ */

static const uint64_t ORC_MULTIPLE = 6364136223846793005ULL;           /* fallback_hash.rs MULTIPLE */
static const uint64_t ORC_PI[4]  = { 0x243f6a8885a308d3ULL, 0x13198a2e03707344ULL,
                                     0xa4093822299f31d0ULL, 0x082efa98ec4e6c89ULL };
static const uint64_t ORC_PI2[4] = { 0x452821e638d01377ULL, 0xbe5466cf34e90c6cULL,
                                     0xc0ac29b7c97c50ddULL, 0x3f84d5b5b5470917ULL };

/* ahash operations.rs folded_multiply: 128-bit product, low ^ high. */
static inline uint64_t folded_multiply(uint64_t s, uint64_t by) {
    unsigned __int128 r = (unsigned __int128)s * (unsigned __int128)by;
    return (uint64_t)r ^ (uint64_t)(r >> 64);
}
static inline uint64_t rotl64(uint64_t x, unsigned r) { r &= 63; return r ? (x << r) | (x >> (64 - r)) : x; }

/* Full fallback AHasher (buffer, pad): update = write_u64; finish rotates by buffer&63. */
typedef struct { uint64_t buffer, pad; } ahasher;
static inline void ah_write_u64(ahasher *h, uint64_t v) { h->buffer = folded_multiply(v ^ h->buffer, ORC_MULTIPLE); }
static inline uint64_t ah_finish(const ahasher *h) {
    return rotl64(folded_multiply(h->buffer, h->pad), (unsigned)(h->buffer & 63));
}

/* RandomState::from_keys(a, b, c) (ahash random_state.rs): hasher from a (buffer=k1, pad=k0),
 * write_usize(c), k_i = mix(b..) with mix(l,r) = clone.write_u64(l).write_u64(r).finish(). */
static void orc_from_keys(const uint64_t a[4], const uint64_t b[4], uint64_t c, uint64_t out[4]) {
    ahasher base = { a[1], a[0] };
    ah_write_u64(&base, c);
    const int l[4] = {0, 1, 2, 3}, r[4] = {2, 3, 1, 0};
    for (int i = 0; i < 4; i++) {
        ahasher h = base;
        ah_write_u64(&h, b[l[i]]);
        ah_write_u64(&h, b[r[i]]);
        out[i] = ah_finish(&h);
    }
}

/* RandomState::with_seed(seed) with get_fixed_seeds() == [PI, PI2] (no rng feature).
 * Call site: src/shared/shared.rs:14 `RandomState::with_seed(0)`. */
void orc_keys_with_seed(uint64_t seed, uint64_t out[4]) { orc_from_keys(ORC_PI, ORC_PI2, seed, out); }

/* hash_one for a 64-bit-or-narrower integer.
 *   flags bit0 = 0: `specialize` path (nightly; the reference is nightly-only): AHasherU64
 *                   {buffer:k1, pad:k0}; write_u64; finish = folded_multiply(buffer, pad)
 *   flags bit0 = 1: stable path: full AHasher, finish additionally rotates by buffer & 63.
 * i32/u32 keys arrive zero-extended from their 32-bit pattern, i64 reinterpret as u64. */
uint64_t orc_hash_u64(const uint64_t k[4], uint64_t v, int flags) {
    uint64_t buffer = folded_multiply(v ^ k[1], ORC_MULTIPLE);
    uint64_t h = folded_multiply(buffer, k[0]);
    return (flags & 1) ? rotl64(h, (unsigned)(buffer & 63)) : h;
}

/* hash_one for a Utf8 value (datafusion-common 40 hash_utils: `impl HashValue for str` -> state.hash_one(&str)).
 * [recalled, unpinned] std's `Hash for str` is write(bytes) then write_u8(0xff); ahash 0.8.11 fallback_hash.rs:
 *   write(bytes): buffer = (buffer + len) * MULTIPLE; then
 *       len <= 8 : large_update(read_small(bytes));  9..16: large_update([first u64, last u64]);
 *       len > 16 : large_update(last 16 bytes), then large_update(every leading 16-byte block while > 16 bytes remain)
 *   large_update([a, b]): buffer = rotl((buffer + pad) ^ folded_multiply(a ^ extra[0], b ^ extra[1]), 23)
 *   write_u8(x) = update(x): buffer = folded_multiply(x ^ buffer, MULTIPLE);  finish rotates by buffer & 63.
 * AHasher::from_random_state: buffer = k1, pad = k0, extra = [k2, k3]. Join RESULTS do not depend on this value
 * (equal_rows_arr compares the strings); the GPU hash is checked bit for bit against this function. */
static inline uint64_t rd_le(const uint8_t *p, int n) { uint64_t v = 0; memcpy(&v, p, (size_t)n); return v; }
uint64_t orc_hash_bytes(const uint64_t k[4], const uint8_t *d, uint64_t len) {
    uint64_t buffer = k[1];
    const uint64_t pad = k[0];
    buffer = (buffer + len) * ORC_MULTIPLE;
#define ORC_LARGE(a, b) buffer = rotl64((buffer + pad) ^ folded_multiply((a) ^ k[2], (b) ^ k[3]), 23)
    if (len > 8) {
        if (len > 16) {
            ORC_LARGE(rd_le(d + len - 16, 8), rd_le(d + len - 8, 8));
            uint64_t off = 0;
            while (len - off > 16) { ORC_LARGE(rd_le(d + off, 8), rd_le(d + off + 8, 8)); off += 16; }
        } else {
            ORC_LARGE(rd_le(d, 8), rd_le(d + len - 8, 8));
        }
    } else {
        uint64_t a = 0, b = 0;                                    /* read_small */
        if (len >= 4) { a = rd_le(d, 4); b = rd_le(d + len - 4, 4); }
        else if (len >= 2) { a = rd_le(d, 2); b = d[len - 1]; }
        else if (len == 1) { a = d[0]; b = d[0]; }
        ORC_LARGE(a, b);
    }
#undef ORC_LARGE
    buffer = folded_multiply(0xffULL ^ buffer, ORC_MULTIPLE);      /* write_u8(0xff) of Hash for str */
    return rotl64(folded_multiply(buffer, pad), (unsigned)(buffer & 63));
}

/* datafusion-common hash_utils::combine_hashes. */
static inline uint64_t combine_hashes(uint64_t l, uint64_t r) {
    uint64_t h = (uint64_t)(17 * 37) + l;
    return h * 37 + r;
}

static inline int col_valid(const orc_col *c, size_t i) {
    return !c->validity || ((c->validity[i >> 3] >> (i & 7)) & 1);
}
static inline uint64_t col_bits(const orc_col *c, size_t i) {
    if (c->width == 8) return ((const uint64_t *)c->values)[i];
    return (uint64_t)((const uint32_t *)c->values)[i];            /* zero-extended 32-bit pattern */
}

/* calculate_hash (src/shared/shared.rs:11-16) -> create_hashes(cols, with_seed(0), buf):
 * buffer starts at 0; column 0: hash = hash_one(v); column j>=1: hash = combine_hashes(hash_one(v), hash);
 * a NULL slot leaves the running hash unchanged. */
void orc_create_hashes(int32_t n_cols, const orc_col *cols, uint64_t n_rows,
                       const uint64_t k[4], int32_t flags, uint64_t *out) {
    for (uint64_t i = 0; i < n_rows; i++) out[i] = 0;
    for (int32_t c = 0; c < n_cols; c++) {
        for (uint64_t i = 0; i < n_rows; i++) {
            if (!col_valid(&cols[c], i)) continue;
            uint64_t h = cols[c].width == 0
                ? orc_hash_bytes(k, (const uint8_t *)cols[c].values + cols[c].offsets[i], (uint64_t)(cols[c].offsets[i + 1] - cols[c].offsets[i]))
                : orc_hash_u64(k, col_bits(&cols[c], i), flags);
            out[i] = (c == 0) ? h : combine_hashes(h, out[i]);
        }
    }
}

/* ------------------------------------------------------------------------------------------ */
/* Build side: hash -> head row (+1), LIFO overflow chain (version10 semantics)               */
/* ------------------------------------------------------------------------------------------ */

typedef struct {
    uint64_t *slot_hash;    /* 0 = empty, else hash | 1<<63 (fixed_table.rs:618) */
    uint64_t *slot_head;    /* head row + 1 */
    uint64_t  cap;          /* power of two */
    uint64_t  used;
    uint64_t *overflow;     /* overflow[row+1] = previous head (+1) or 0 */
    uint64_t  rows;         /* global row counter == offset_counter (parallel_join_execution_state.rs:95) */
    uint64_t  overflow_cap;
} orc_map;

orc_map *orc_map_new(void) {
    orc_map *m = (orc_map *)calloc(1, sizeof(orc_map));
    m->cap = 16;                                              /* new_map_3.rs:162 starts at 16 slots */
    m->slot_hash = (uint64_t *)calloc(m->cap, 8);
    m->slot_head = (uint64_t *)calloc(m->cap, 8);
    m->overflow_cap = 16;
    m->overflow = (uint64_t *)calloc(m->overflow_cap, 8);
    return m;
}
void orc_map_free(orc_map *m) {
    if (!m) return;
    free(m->slot_hash); free(m->slot_head); free(m->overflow); free(m);
}
uint64_t orc_map_rows(const orc_map *m) { return m->rows; }

static uint64_t map_insert_raw(orc_map *m, uint64_t stored, uint64_t head) {
    uint64_t mask = m->cap - 1, i = stored & mask;
    for (;;) {
        if (m->slot_hash[i] == 0) { m->slot_hash[i] = stored; m->slot_head[i] = head; m->used++; return 0; }
        if (m->slot_hash[i] == stored) { uint64_t p = m->slot_head[i]; m->slot_head[i] = head; return p; }
        i = (i + 1) & mask;
    }
}
static void map_grow(orc_map *m) {                            /* new_map_3.rs:325-411 doubles the table */
    uint64_t ocap = m->cap, *oh = m->slot_hash, *ov = m->slot_head;
    m->cap = ocap * 2; m->used = 0;
    m->slot_hash = (uint64_t *)calloc(m->cap, 8);
    m->slot_head = (uint64_t *)calloc(m->cap, 8);
    for (uint64_t i = 0; i < ocap; i++) if (oh[i]) map_insert_raw(m, oh[i], ov[i]);
    free(oh); free(ov);
}

/* JoinStateInstance::add (version10/parallel_join_execution_state.rs:91-133):
 * offset = fetch_add(n); prev = table.insert(hash, offset+i+1); overflow[offset+i+1] = prev. */
void orc_map_add(orc_map *m, const uint64_t *hashes, uint64_t n) {
    uint64_t offset = m->rows;
    if (offset + n + 1 > m->overflow_cap) {
        uint64_t ncap = m->overflow_cap;
        while (ncap < offset + n + 1) ncap *= 2;
        m->overflow = (uint64_t *)realloc(m->overflow, ncap * 8);
        memset(m->overflow + m->overflow_cap, 0, (ncap - m->overflow_cap) * 8);
        m->overflow_cap = ncap;
    }
    for (uint64_t i = 0; i < n; i++) {
        if ((m->used + 1) * 8 > m->cap * 7) map_grow(m);     /* capacity rule fixed_table.rs:1008-1015 */
        uint64_t prev = map_insert_raw(m, hashes[i] | (1ULL << 63), offset + i + 1);
        m->overflow[offset + i + 1] = prev;
    }
    m->rows = offset + n;
}

/* Version10Lookup::get_iter (lookup_implementation_3.rs:31-59): idx = table.get(hash);
 * while idx != 0 { yield idx-1; idx = overflow[idx]; }  — most-recent-first. */
uint64_t orc_map_get_all(const orc_map *m, uint64_t hash, uint64_t *out, uint64_t cap) {
    uint64_t stored = hash | (1ULL << 63), mask = m->cap - 1, i = stored & mask, n = 0;
    for (;;) {
        if (m->slot_hash[i] == 0) return 0;
        if (m->slot_hash[i] == stored) break;
        i = (i + 1) & mask;
    }
    for (uint64_t idx = m->slot_head[i]; idx != 0; idx = m->overflow[idx]) {
        if (n < cap) out[n] = idx - 1;
        n++;
    }
    return n;
}

/* ------------------------------------------------------------------------------------------ */
/* ConcurrentBitSet (src/utils/concurrent_bit_set.rs)                                          */
/* ------------------------------------------------------------------------------------------ */

typedef struct { uint64_t length, set_count, words; uint64_t *contents; } orc_bitset;

orc_bitset *orc_bitset_new(uint64_t length) {                  /* with_capacity :17-26 */
    orc_bitset *b = (orc_bitset *)calloc(1, sizeof(orc_bitset));
    b->length = length; b->words = (length + 63) / 64;
    b->contents = (uint64_t *)calloc(b->words ? b->words : 1, 8);
    return b;
}
void orc_bitset_free(orc_bitset *b) { if (b) { free(b->contents); free(b); } }
/* set_ones :28-52 + write_bits_at :54-60 (semantics: OR the bits in, count newly set). */
int32_t orc_bitset_set_ones(orc_bitset *b, const uint64_t *idx, uint64_t n) {
    for (uint64_t i = 0; i < n; i++) {
        if (idx[i] >= b->length) return -1;                    /* "Index out of bounds" panic :31 */
        uint64_t bit = 1ULL << (idx[i] & 63), *w = &b->contents[idx[i] >> 6];
        if (!(*w & bit)) { *w |= bit; b->set_count++; }
    }
    return 0;
}
uint64_t orc_bitset_set_count(const orc_bitset *b) { return b->set_count; }
/* get_unset_indices_array :117-131; get_set_indices_array :93-106 with the :99 bug NOT reproduced
 * (`bits & (1<<i) == 1` only holds for i == 0; the intended semantics is `!= 0`). */
uint64_t orc_bitset_indices(const orc_bitset *b, int32_t want_set, uint64_t *out) {
    uint64_t n = 0;
    for (uint64_t i = 0; i < b->length; i++) {
        int set = (int)((b->contents[i >> 6] >> (i & 63)) & 1);
        if (set == (want_set != 0)) out[n++] = i;
    }
    return n;
}
/* get_set_indices :85-91 as a byte-per-row boolean vector. */
void orc_bitset_bools(const orc_bitset *b, uint8_t *out) {
    for (uint64_t i = 0; i < b->length; i++) out[i] = (uint8_t)((b->contents[i >> 6] >> (i & 63)) & 1);
}

/* ------------------------------------------------------------------------------------------ */
/* Index-pair vectors                                                                         */
/* ------------------------------------------------------------------------------------------ */

typedef struct { uint64_t *build; uint32_t *probe; uint64_t n, cap; } orc_pairs;

orc_pairs *orc_pairs_new(void) { return (orc_pairs *)calloc(1, sizeof(orc_pairs)); }
void orc_pairs_free(orc_pairs *p) { if (p) { free(p->build); free(p->probe); free(p); } }
uint64_t orc_pairs_len(const orc_pairs *p) { return p->n; }
const uint64_t *orc_pairs_build(const orc_pairs *p) { return p->build; }
const uint32_t *orc_pairs_probe(const orc_pairs *p) { return p->probe; }
void orc_pairs_clear(orc_pairs *p) { p->n = 0; }
void orc_pairs_push(orc_pairs *p, uint64_t b, uint32_t q) {
    if (p->n == p->cap) {
        p->cap = p->cap ? p->cap * 2 : 1024;
        p->build = (uint64_t *)realloc(p->build, p->cap * 8);
        p->probe = (uint32_t *)realloc(p->probe, p->cap * 4);
    }
    p->build[p->n] = b; p->probe[p->n] = q; p->n++;
}

/* ------------------------------------------------------------------------------------------ */
/* Probe: hash -> candidates -> key equality                                                  */
/* ------------------------------------------------------------------------------------------ */

/* get_matching_indices (shared.rs:29-47) when with_probe == 0,
 * get_matching_indices_with_probe (shared.rs:49-81) when with_probe != 0 (emits (probe, NULL) for misses). */
void orc_get_matching_indices(const orc_map *m, const uint64_t *probe_hashes, uint64_t n_probe,
                              int32_t with_probe, orc_pairs *out) {
    out->n = 0;
    for (uint64_t p = 0; p < n_probe; p++) {
        uint64_t stored = probe_hashes[p] | (1ULL << 63), mask = m->cap - 1, i = stored & mask, cnt = 0;
        uint64_t head = 0;
        for (;;) {
            if (m->slot_hash[i] == 0) break;
            if (m->slot_hash[i] == stored) { head = m->slot_head[i]; break; }
            i = (i + 1) & mask;
        }
        for (uint64_t idx = head; idx != 0; idx = m->overflow[idx]) { orc_pairs_push(out, idx - 1, (uint32_t)p); cnt++; }
        if (with_probe && cnt == 0) orc_pairs_push(out, ORC_NULL64, (uint32_t)p);
    }
}

/* equal_rows_arr (datafusion_private.rs:40-80) with null_equals_null == false: keep a pair iff every
 * key column is valid on both sides and the values are equal; a NULL build index compares as NULL. */
void orc_equal_rows(const orc_pairs *in, int32_t n_cols, const orc_col *build_cols, const orc_col *probe_cols,
                    orc_pairs *out) {
    out->n = 0;
    for (uint64_t j = 0; j < in->n; j++) {
        uint64_t b = in->build[j]; uint32_t p = in->probe[j];
        if (b == ORC_NULL64) continue;
        int eq = 1;
        for (int32_t c = 0; c < n_cols && eq; c++) {
            if (!col_valid(&build_cols[c], b) || !col_valid(&probe_cols[c], p)) { eq = 0; break; }
            if (build_cols[c].width != probe_cols[c].width) { eq = 0; break; }
            if (build_cols[c].width == 0) {
                const int32_t bo = build_cols[c].offsets[b], bl = build_cols[c].offsets[b + 1] - bo;
                const int32_t po = probe_cols[c].offsets[p], pl = probe_cols[c].offsets[p + 1] - po;
                if (bl != pl || memcmp((const char *)build_cols[c].values + bo, (const char *)probe_cols[c].values + po, (size_t)bl) != 0) eq = 0;
            } else if (col_bits(&build_cols[c], b) != col_bits(&probe_cols[c], p)) eq = 0;
        }
        if (eq) orc_pairs_push(out, b, p);
    }
}

/* Steps 1-3 of every lookup_*_probe_batch (e.g. inner.rs:92-108, full.rs:128-145):
 * hash the probe keys, enumerate candidates, drop hash collisions and NULL keys. */
void orc_match_pairs(const orc_map *m, int32_t join_type, int32_t n_cols,
                     const orc_col *build_cols, const orc_col *probe_cols, uint64_t n_probe,
                     const uint64_t k[4], int32_t hash_flags, orc_pairs *out) {
    uint64_t *hashes = (uint64_t *)malloc((n_probe ? n_probe : 1) * 8);
    orc_create_hashes(n_cols, probe_cols, n_probe, k, hash_flags, hashes);
    orc_pairs *cand = orc_pairs_new();
    int with_probe = (join_type == ORC_RIGHT || join_type == ORC_FULL);   /* right_outer.rs:90-92, full.rs:132-133 */
    orc_get_matching_indices(m, hashes, n_probe, with_probe, cand);
    orc_equal_rows(cand, n_cols, build_cols, probe_cols, out);
    orc_pairs_free(cand); free(hashes);
}

/* Join-type post-processing of one probe batch, applied to the (possibly JoinFilter-ed) matched pairs:
 *   Inner      inner.rs:124-128          pairs as they are
 *   Left       left_outer.rs:158-168     mark visited; pairs as they are
 *   Right      right_outer.rs:120-125    append_right_indices(.., 0..n_probe, false)  datafusion_private.rs:157-185
 *   Full       full.rs:160-171           mark visited + append_right_indices
 *   LeftSemi   left_semi.rs:158-163      mark visited; no per-batch output
 *   LeftAnti   left_anti.rs:158-163      mark visited; no per-batch output
 *   RightSemi  right_semi.rs:120-121     get_semi_indices (datafusion_private.rs:114-140): unique matched probe rows
 *   RightAnti  right_anti.rs:121-122     get_anti_indices (datafusion_private.rs:85-111): unmatched probe rows
 * Output NULLs are ORC_NULL64 / ORC_NULL32. */
int32_t orc_finish_batch(int32_t join_type, const orc_pairs *matched, uint64_t n_probe,
                         orc_bitset *visited, orc_pairs *out) {
    out->n = 0;
    int mark = (join_type == ORC_LEFT || join_type == ORC_FULL ||
                join_type == ORC_LEFT_SEMI || join_type == ORC_LEFT_ANTI);
    if (mark) {
        if (!visited) return -1;
        for (uint64_t j = 0; j < matched->n; j++)
            if (orc_bitset_set_ones(visited, &matched->build[j], 1) != 0) return -1;
    }
    if (join_type == ORC_LEFT_SEMI || join_type == ORC_LEFT_ANTI) return 0;
    if (join_type == ORC_INNER || join_type == ORC_LEFT || join_type == ORC_RIGHT || join_type == ORC_FULL)
        for (uint64_t j = 0; j < matched->n; j++) orc_pairs_push(out, matched->build[j], matched->probe[j]);
    if (join_type == ORC_INNER || join_type == ORC_LEFT) return 0;

    uint8_t *seen = (uint8_t *)calloc(n_probe ? n_probe : 1, 1);
    for (uint64_t j = 0; j < matched->n; j++) if (matched->probe[j] < n_probe) seen[matched->probe[j]] = 1;
    if (join_type == ORC_RIGHT || join_type == ORC_FULL) {
        for (uint64_t p = 0; p < n_probe; p++) if (!seen[p]) orc_pairs_push(out, ORC_NULL64, (uint32_t)p);
    } else if (join_type == ORC_RIGHT_SEMI) {
        for (uint64_t p = 0; p < n_probe; p++) if (seen[p]) orc_pairs_push(out, ORC_NULL64, (uint32_t)p);
    } else if (join_type == ORC_RIGHT_ANTI) {
        for (uint64_t p = 0; p < n_probe; p++) if (!seen[p]) orc_pairs_push(out, ORC_NULL64, (uint32_t)p);
    }
    free(seen);
    return 0;
}

/* Finaliser batch, emitted once by the last partition stream:
 *   Left/Full  emit_unmatched_build_records  left_outer.rs:174-195, full.rs:181-201  (unset rows, probe side NULL)
 *   LeftSemi   emit_matched_build_records    left_semi.rs:166-179                    (set rows)
 *   LeftAnti   left_anti.rs:166-179                                                  (unset rows)
 * Returns build-row indices with probe == ORC_NULL32. */
int32_t orc_finish_join(int32_t join_type, const orc_bitset *visited, orc_pairs *out) {
    out->n = 0;
    if (join_type == ORC_INNER || join_type == ORC_RIGHT || join_type == ORC_RIGHT_SEMI || join_type == ORC_RIGHT_ANTI)
        return 0;
    if (!visited) return -1;
    int want_set = (join_type == ORC_LEFT_SEMI);
    for (uint64_t i = 0; i < visited->length; i++) {
        int set = (int)((visited->contents[i >> 6] >> (i & 63)) & 1);
        if (set == want_set) orc_pairs_push(out, i, ORC_NULL32);
    }
    return 0;
}

/* ------------------------------------------------------------------------------------------ */
/* Workload generators (restated for the measurement inputs)                    */
/* ------------------------------------------------------------------------------------------ */

/* make_exponential_int_array (src/api_utils.rs:15-23), all arithmetic in f32, truncating cast.
 * KAT (api_utils.rs:51-71): (0,10) -> [0,0,0,0,1,2,2,3,5,7]. */
void orc_make_exponential_int_array(int32_t min, int32_t max, int32_t *out) {
    const float base = 16.0f;
    int32_t diff = max - min;
    for (int32_t n = 0; n < diff; n++) {
        float x = (float)n / (float)diff;
        float y = (powf(base, x) - 1.0f) / (base - 1.0f);
        out[n] = min + (int32_t)(y * (float)diff);
    }
}

/* make_int_array_with_shift (src/api_utils.rs:6-9). */
void orc_make_int_array_with_shift(int32_t min, int32_t max, int32_t shift, int32_t *out) {
    for (int32_t v = min + shift, i = 0; v < max + shift; v++, i++) out[i] = v;
}

static inline uint64_t splitmix64(uint64_t *s) {
    uint64_t z = (*s += 0x9e3779b97f4a7c15ULL);
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
    return z ^ (z >> 31);
}
/* Random permutation of 0..n-1 (Fisher-Yates over splitmix64(seed)). */
void orc_gen_permutation_i64(uint64_t n, uint64_t seed, int64_t *out) {
    for (uint64_t i = 0; i < n; i++) out[i] = (int64_t)i;
    uint64_t s = seed;
    for (uint64_t i = n; i > 1; i--) {
        uint64_t j = splitmix64(&s) % i;
        int64_t t = out[i - 1]; out[i - 1] = out[j]; out[j] = t;
    }
}
/* Uniform int64 in [0, range). */
void orc_gen_uniform_i64(uint64_t n, uint64_t range, uint64_t seed, int64_t *out) {
    uint64_t s = seed;
    for (uint64_t i = 0; i < n; i++) out[i] = (int64_t)(splitmix64(&s) % range);
}
