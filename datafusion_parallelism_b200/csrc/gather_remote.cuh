// gather_remote.cuh — late materialisation of join output ROWS when the sides were handed over as PIECES (the multi-GPU exchange:
// one piece per source rank). take_multiple_record_batch (src/shared/shared.rs:83-92) across GPUs: the local join's
// indices count the rows this rank RECEIVED; the column values stayed on the rank that owns the row. Three steps, all on the
// consuming GPU, every source read through plain (cudaIpc-mapped peer) pointers over NVLink:
//
//   resolve_origin_kernel      local index -> (piece, row id at the piece's source)      one coalesced index read, one 4-byte
//                                                                                         read of the piece's row-id array
//   gather_pieces_kernel       (piece, row) -> column value + validity word              one `width`-byte read per row from the
//                                                                                         source rank's column
//   gather_pieces_utf8_*       (piece, row) -> Arrow offsets, then the bytes             lengths, scan (gather.cuh), warp-per-row copy
//
// Reads over NVLink are latency-bound like the local gather: a warp takes 128 output rows, lane l the rows l, l+32, l+64, l+96,
// so every thread has four independent remote reads in flight and the ballot of an item IS the output validity word.
#pragma once
#include "gather.cuh"

namespace b200j {

constexpr int GP_MAX_PIECES = 64;
constexpr int GP_ITEMS = 4;
constexpr uint32_t GP_NULL_ROW = 0xffffffffu;

struct OriginPieces {
    const uint32_t *rowid[GP_MAX_PIECES];      // row ids of the piece's rows at their source (device or peer memory); null = the position itself
    uint32_t        end[GP_MAX_PIECES];        // exclusive end of the piece in the concatenated (received) batch
    uint32_t        rowid_base[GP_MAX_PIECES]; // added to what rowid[] holds (identity pieces: first row of the piece at its source)
    int32_t         n;
};

// piece holding position x: pieces are few (<= 64) and the table sits in constant kernel-parameter space; branch-free binary search
__device__ __forceinline__ uint32_t piece_of(const OriginPieces &op, uint32_t x) {
    uint32_t lo = 0, n = (uint32_t)op.n;
    while (n > 1) {
        const uint32_t half = n >> 1;
        lo = (op.end[lo + half - 1] <= x) ? lo + half : lo;
        n -= half;
    }
    return lo;
}

template <typename IdxT>
__global__ void __launch_bounds__(256) resolve_origin_kernel(OriginPieces op, const IdxT *__restrict__ idx, int64_t n, uint32_t total,
                                                             uint32_t *__restrict__ out_piece, uint32_t *__restrict__ out_row) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t row0 = (((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5) * (32 * GP_ITEMS); row0 < n; row0 += warps * (32 * GP_ITEMS)) {
        uint32_t pc[GP_ITEMS], rr[GP_ITEMS];
#pragma unroll
        for (int q = 0; q < GP_ITEMS; q++) {
            const int64_t i = row0 + q * 32 + lane;
            pc[q] = 0;
            rr[q] = GP_NULL_ROW;
            if (i < n) {
                const IdxT x = __ldcs(idx + i);
                if (!idx_is_null(x) && (unsigned long long)x < (unsigned long long)total) {
                    const uint32_t p = piece_of(op, (uint32_t)x);
                    const uint32_t start = p ? op.end[p - 1] : 0u;
                    const uint32_t k = (uint32_t)x - start;
                    pc[q] = p;
                    rr[q] = (op.rowid[p] ? __ldg(op.rowid[p] + k) : k) + op.rowid_base[p];
                }
            }
        }
#pragma unroll
        for (int q = 0; q < GP_ITEMS; q++) {
            const int64_t i = row0 + q * 32 + lane;
            if (i < n) {
                out_piece[i] = pc[q];
                out_row[i] = rr[q];
            }
        }
    }
}

struct PieceCols {
    const void    *values[GP_MAX_PIECES];      // the column on the piece's source rank (fixed width: values; Utf8: int32 offsets)
    const uint8_t *validity[GP_MAX_PIECES];    // LSB bitmap starting at the column's row 0, or null
    const uint8_t *data[GP_MAX_PIECES];        // Utf8: the bytes
    int32_t        n;
};

template <typename T>
__global__ void __launch_bounds__(256) gather_pieces_kernel(PieceCols pc, const uint32_t *__restrict__ piece, const uint32_t *__restrict__ row, int64_t n,
                                                            T *__restrict__ dst, uint32_t *__restrict__ dst_valid, unsigned long long *null_count) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    uint32_t nulls = 0;
    for (int64_t row0 = (((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5) * (32 * GP_ITEMS); row0 < n; row0 += warps * (32 * GP_ITEMS)) {
        T v[GP_ITEMS];
        bool valid[GP_ITEMS];
#pragma unroll
        for (int q = 0; q < GP_ITEMS; q++) {
            const int64_t i = row0 + q * 32 + lane;
            v[q] = T{};
            valid[q] = false;
            if (i < n) {
                const uint32_t r = __ldcs(row + i);
                const uint32_t p = __ldcs(piece + i);
                if (r != GP_NULL_ROW && p < (uint32_t)pc.n && pc.values[p] != nullptr && bit_valid(pc.validity[p], (int64_t)r)) {
                    v[q] = reinterpret_cast<const T *>(pc.values[p])[r];
                    valid[q] = true;
                }
            }
        }
#pragma unroll
        for (int q = 0; q < GP_ITEMS; q++) {
            const int64_t i = row0 + q * 32 + lane;
            if (i < n) {
                dst[i] = v[q];
                if (!valid[q]) nulls++;
            }
            const uint32_t word = __ballot_sync(0xffffffffu, valid[q]);
            if (lane == 0 && row0 + q * 32 < n && dst_valid) dst_valid[(row0 + q * 32) >> 5] = word;
        }
    }
    const unsigned long long z = warp_sum_u32(nulls);
    if (lane == 0 && z && null_count) atomicAdd(null_count, z);
}

// Utf8 pass 1: byte length of every output row (0 for NULL) + validity word + null count
__global__ void __launch_bounds__(256) gather_pieces_utf8_len_kernel(PieceCols pc, const uint32_t *__restrict__ piece, const uint32_t *__restrict__ row,
                                                                     int64_t n, int64_t *__restrict__ len, uint32_t *__restrict__ dst_valid,
                                                                     unsigned long long *null_count) {
    const int lane = threadIdx.x & 31;
    uint32_t nulls = 0;
    const int64_t n_round = (n + 31) & ~(int64_t)31;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += (int64_t)gridDim.x * blockDim.x) {
        bool valid = false;
        if (i < n) {
            const uint32_t r = row[i], p = piece[i];
            int64_t l = 0;
            if (r != GP_NULL_ROW && p < (uint32_t)pc.n && pc.values[p] != nullptr && bit_valid(pc.validity[p], (int64_t)r)) {
                const int32_t *off = reinterpret_cast<const int32_t *>(pc.values[p]);
                l = (int64_t)off[r + 1] - (int64_t)off[r];
                valid = true;
            }
            len[i] = l;
            if (!valid) nulls++;
        }
        const uint32_t word = __ballot_sync(0xffffffffu, valid);
        if (lane == 0 && dst_valid) dst_valid[i >> 5] = word;
    }
    const unsigned long long z = warp_sum_u32(nulls);
    if (lane == 0 && z && null_count) atomicAdd(null_count, z);
}

// Utf8 pass 2: one warp per output row copies its bytes from the source rank
__global__ void __launch_bounds__(256) gather_pieces_utf8_copy_kernel(PieceCols pc, const uint32_t *__restrict__ piece, const uint32_t *__restrict__ row,
                                                                      int64_t n, const int32_t *__restrict__ out_off, uint8_t *__restrict__ dst) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < n; i += warps) {
        const int64_t d = out_off[i], l = (int64_t)out_off[i + 1] - d;
        if (l <= 0) continue;
        const uint32_t r = row[i], p = piece[i];
        const int32_t *off = reinterpret_cast<const int32_t *>(pc.values[p]);
        const uint8_t *sp = pc.data[p] + off[r];
        for (int64_t k = lane; k < l; k += 32) dst[d + k] = sp[k];
    }
}

// out[i] = 1 when row i of two Arrow Utf8 columns holds the same bytes (the key check of a join keyed on row hashes; NULL rows are
// never equal: equal_rows_arr with null_equals_null = false, src/shared/datafusion_private.rs:40-80)
__global__ void __launch_bounds__(256) utf8_rows_equal_kernel(const int32_t *__restrict__ off_a, const uint8_t *__restrict__ data_a,
                                                              const uint32_t *__restrict__ valid_a, const int32_t *__restrict__ off_b,
                                                              const uint8_t *__restrict__ data_b, const uint32_t *__restrict__ valid_b, int64_t n,
                                                              uint8_t *__restrict__ out) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        bool eq = (!valid_a || ((valid_a[i >> 5] >> (i & 31)) & 1u)) && (!valid_b || ((valid_b[i >> 5] >> (i & 31)) & 1u));
        const int32_t a0 = off_a[i], b0 = off_b[i];
        const int32_t la = off_a[i + 1] - a0, lb = off_b[i + 1] - b0;
        eq = eq && la == lb;
        for (int32_t k = 0; eq && k < la; k++) eq = data_a[a0 + k] == data_b[b0 + k];
        out[i] = eq ? 1 : 0;
    }
}

}  // namespace b200j
