/*
 * oracle/cpu_join_mt.c — multi-threaded CPU stand-in for "the reference's fastest CPU join".
 *
 * TEST / MEASUREMENT INFRASTRUCTURE ONLY (bench.py cpu_baseline and --impl reference legs).
 * The Rust original cannot be built on this box (no cargo; nightly; version10 is aarch64-only), so
 * this is a C restatement ("port") of version10's ALGORITHM, run with all host threads:
 *
 *   build  JoinStateInstance::add      version10/parallel_join_execution_state.rs:91-133
 *          Inner::insert_atomically    version10/new_map_3/fixed_table.rs:560-672
 *            lock-free open addressing keyed by the 64-bit row hash: compare_exchange(0 -> hash|1<<63)
 *            on the slot's hash word (:618), swap() of the head word (:591,:623), overflow[row+1] = prev
 *   probe  get_matching_indices        src/shared/shared.rs:29-47 (per 8192-row batch)
 *          Version10Lookup::get_iter   version10/lookup_implementation_3.rs:31-59 (LIFO chain walk)
 *          equal_rows_arr              src/shared/datafusion_private.rs:40-80 (key equality on candidates)
 *   hash   calculate_hash              src/shared/shared.rs:11-16 (ahash fallback, see join_oracle.c)
 *
 * Deliberate simplifications, all in the baseline's favour: the table is pre-sized once to
 * next_pow2(n*8/7) (the reference starts at 16 slots, doubles and re-inserts everything at finish,
 * new_map_3.rs:162,325-411,182-247); linear probing replaces the NEON tag groups; the build side is
 * not concat-copied (parallel_join_execution_state.rs:256-298).
 */
#define _GNU_SOURCE
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <stdatomic.h>
#include <omp.h>

uint64_t orc_hash_u64(const uint64_t k[4], uint64_t v, int flags);   /* join_oracle.c */

#define BATCH_ROWS 8192   /* DataFusion default batch size; the benches use 8192-row batches (build_speed.rs:53-64) */

typedef struct {
    _Atomic uint64_t *slot_hash;
    _Atomic uint64_t *slot_head;
    uint64_t          cap;
    uint64_t         *overflow;     /* [n_build + 1] */
    const int64_t    *build_keys;
    uint64_t          n_build;
} mt_table;

static double now_s(void) {
    struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

mt_table *mt_table_new(uint64_t n_build) {
    mt_table *t = (mt_table *)calloc(1, sizeof(mt_table));
    uint64_t want = n_build * 8 / 7 + 1, cap = 16;
    while (cap < want) cap <<= 1;
    t->cap = cap; t->n_build = n_build;
    t->slot_hash = (_Atomic uint64_t *)aligned_alloc(64, cap * 8);
    t->slot_head = (_Atomic uint64_t *)aligned_alloc(64, cap * 8);
    t->overflow = (uint64_t *)aligned_alloc(64, ((n_build + 1) * 8 + 63) / 64 * 64);
    return t;
}
void mt_table_free(mt_table *t) {
    if (!t) return;
    free((void *)t->slot_hash); free((void *)t->slot_head); free(t->overflow); free(t);
}

/* Timed build: returns seconds.  threads <= 0 -> omp default. */
double mt_build(mt_table *t, const int64_t *keys, uint64_t n, const uint64_t k[4], int32_t threads) {
    if (threads > 0) omp_set_num_threads(threads);
    t->build_keys = keys;
    const uint64_t mask = t->cap - 1;
    /* zeroing the table is part of the build (the reference allocates zeroed tables) */
    double t0 = now_s();
#pragma omp parallel for schedule(static)
    for (uint64_t i = 0; i < t->cap; i++) {
        atomic_store_explicit(&t->slot_hash[i], 0, memory_order_relaxed);
        atomic_store_explicit(&t->slot_head[i], 0, memory_order_relaxed);
    }
    uint64_t n_batches = (n + BATCH_ROWS - 1) / BATCH_ROWS;
#pragma omp parallel for schedule(dynamic, 4)
    for (uint64_t b = 0; b < n_batches; b++) {
        uint64_t lo = b * BATCH_ROWS, hi = lo + BATCH_ROWS < n ? lo + BATCH_ROWS : n;
        uint64_t hashes[BATCH_ROWS];
        for (uint64_t i = lo; i < hi; i++) hashes[i - lo] = orc_hash_u64(k, (uint64_t)keys[i], 0);
        for (uint64_t i = lo; i < hi; i++) {
            uint64_t stored = hashes[i - lo] | (1ULL << 63), s = stored & mask;
            for (;;) {
                uint64_t cur = atomic_load_explicit(&t->slot_hash[s], memory_order_acquire);
                if (cur == 0) {
                    uint64_t expect = 0;
                    if (atomic_compare_exchange_strong_explicit(&t->slot_hash[s], &expect, stored,
                                                                memory_order_acq_rel, memory_order_acquire))
                        cur = stored;
                    else
                        cur = expect;
                }
                if (cur == stored) {
                    uint64_t prev = atomic_exchange_explicit(&t->slot_head[s], i + 1, memory_order_acq_rel);
                    t->overflow[i + 1] = prev;
                    break;
                }
                s = (s + 1) & mask;
            }
        }
    }
    return now_s() - t0;
}

/* Timed probe: 8192-row batches -> candidate index pairs -> key-equality filter.
 * Writes (build_idx u64, probe_idx u32) pairs per batch into thread-local buffers (as the reference
 * materialises them per batch) and returns the total matched count through *n_matched and an
 * order-independent checksum of the matched pairs through *checksum.  Returns seconds. */
double mt_probe(const mt_table *t, const int64_t *probe_keys, uint64_t n, const uint64_t k[4], int32_t threads,
                uint64_t *n_matched, uint64_t *checksum) {
    if (threads > 0) omp_set_num_threads(threads);
    const uint64_t mask = t->cap - 1;
    uint64_t total = 0, csum = 0;
    uint64_t n_batches = (n + BATCH_ROWS - 1) / BATCH_ROWS;
    double t0 = now_s();
#pragma omp parallel reduction(+ : total, csum)
    {
        uint64_t cap = BATCH_ROWS * 2, *bidx = (uint64_t *)malloc(cap * 8);
        uint32_t *pidx = (uint32_t *)malloc(cap * 4);
        uint64_t hashes[BATCH_ROWS];
#pragma omp for schedule(dynamic, 4)
        for (uint64_t b = 0; b < n_batches; b++) {
            uint64_t lo = b * BATCH_ROWS, hi = lo + BATCH_ROWS < n ? lo + BATCH_ROWS : n, m = 0;
            for (uint64_t i = lo; i < hi; i++) hashes[i - lo] = orc_hash_u64(k, (uint64_t)probe_keys[i], 0);
            /* get_matching_indices */
            for (uint64_t i = lo; i < hi; i++) {
                uint64_t stored = hashes[i - lo] | (1ULL << 63), s = stored & mask, head = 0;
                for (;;) {
                    uint64_t cur = atomic_load_explicit(&t->slot_hash[s], memory_order_relaxed);
                    if (cur == 0) break;
                    if (cur == stored) { head = atomic_load_explicit(&t->slot_head[s], memory_order_relaxed); break; }
                    s = (s + 1) & mask;
                }
                for (uint64_t idx = head; idx != 0; idx = t->overflow[idx]) {
                    if (m == cap) {
                        cap *= 2;
                        bidx = (uint64_t *)realloc(bidx, cap * 8);
                        pidx = (uint32_t *)realloc(pidx, cap * 4);
                    }
                    bidx[m] = idx - 1; pidx[m] = (uint32_t)(i - lo); m++;
                }
            }
            /* equal_rows_arr: compare real key values, compact in place */
            uint64_t w = 0;
            for (uint64_t j = 0; j < m; j++) {
                if (t->build_keys[bidx[j]] == probe_keys[lo + pidx[j]]) {
                    bidx[w] = bidx[j]; pidx[w] = pidx[j]; w++;
                    csum += (bidx[j] + 1) * 0x9e3779b97f4a7c15ULL ^ (lo + pidx[j]);
                }
            }
            total += w;
        }
        free(bidx); free(pidx);
    }
    double dt = now_s() - t0;
    if (n_matched) *n_matched = total;
    if (checksum) *checksum = csum;
    return dt;
}

int32_t mt_max_threads(void) { return omp_get_max_threads(); }
