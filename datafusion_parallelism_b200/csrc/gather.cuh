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

// Utf8 pass 2: one warp per output row copies its bytes
template <typename IdxT>
__global__ void __launch_bounds__(256) gather_utf8_copy_kernel(GatherSrc src, const IdxT *__restrict__ idx, int64_t n, const int64_t *__restrict__ start,
                                                               const int64_t *__restrict__ len, char *__restrict__ dst) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < n; i += warps) {
        const int64_t l = len[i];
        if (l == 0) continue;
        const int64_t s = utf8_off(src, (int64_t)idx[i]), d = start[i];
        const char *sp = reinterpret_cast<const char *>(src.values) + s;
        for (int64_t k = lane; k < l; k += 32) dst[d + k] = sp[k];
    }
}

}  // namespace b200j
