// Micro-benchmark (round 2): what bounds a hash-table probe on B200 when the table is L2-resident per partition?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/ubench/probe_bounds scripts/ubench/probe_bounds.cu
// Every variant does one random 32 B bucket load per row (the hash-table lookup); the variants differ in what streams
// beside it and in how the rows are laid out:
//   G   gather only: keys computed from the row index, result reduced into a register (no streams)
//   S   contiguous stream: 8 B key in, 8 B build index out per row, warp-contiguous 8 B loads / stores
//   P   S with the next iteration's keys prefetched into registers before this iteration's lookups are consumed
//   W   P over (partition, block) SEGMENTS: rows of partition p sit in 128 .. 512-row pieces, one per block of the
//       block-locally partitioned probe stream (round-2 layout); a warp walks one segment at a time
// Table regimes:
//   L2  one 32 MB table (always L2-resident)
//   PT  128 partitions x 30 MB, rows visited in partition order (the radix layout: every table is pulled from HBM once)
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint64_t mix(uint64_t x) { x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33; return x; }

__device__ __forceinline__ void ldb(const void *p, unsigned long long &a, unsigned long long &b, unsigned long long &c, unsigned long long &d) {
    asm volatile("ld.global.L1::no_allocate.L2::evict_last.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(a), "=l"(b), "=l"(c), "=l"(d) : "l"(p));
}

struct Args {
    const char *tab;           // buckets of 32 B
    uint32_t nb;               // buckets per partition (PT) or in total (L2)
    uint32_t n_part;           // 1 = L2 regime
    const unsigned long long *keys;
    unsigned long long *out;
    uint64_t n;
    uint32_t seg;              // W: rows per segment
    uint32_t n_blk;            // W: blocks
    unsigned int *ctr;
};

// bucket address of a key in partition p
__device__ __forceinline__ const char *bucket(const Args &a, uint32_t p, uint64_t k) {
    const uint64_t h = mix(k);
    return a.tab + ((uint64_t)p * a.nb + __umulhi((uint32_t)h, a.nb)) * 32ull;
}

template <int ITEMS, int MODE, int MINB>      // MODE 0 = G, 1 = S, 2 = P
__global__ void __launch_bounds__(256, MINB) stream_kernel(Args a) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr uint32_t WT = 32 * ITEMS;                 // rows per warp iteration
    constexpr uint32_t SUPER = 16384;
    __shared__ unsigned int s_next;
    const uint64_t rows_per_part = a.n / a.n_part;
    unsigned long long acc = 0;
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) s_next = atomicAdd(a.ctr, 1u);
        __syncthreads();
        const uint64_t s0 = (uint64_t)s_next * SUPER;
        if (s0 >= a.n) break;
        const uint32_t p = (uint32_t)(s0 / rows_per_part) < a.n_part ? (uint32_t)(s0 / rows_per_part) : a.n_part - 1;
        unsigned long long key[ITEMS], nxt[ITEMS];
        uint64_t wb = s0 + (uint64_t)warp * WT;
        if (MODE == 2) {
#pragma unroll
            for (int q = 0; q < ITEMS; q++) nxt[q] = __ldcs(a.keys + wb + q * 32 + lane);
        }
        for (; wb < s0 + SUPER; wb += 8 * WT) {
#pragma unroll
            for (int q = 0; q < ITEMS; q++) {
                if (MODE == 0) key[q] = mix(wb + q * 32 + lane);
                else if (MODE == 1) key[q] = __ldcs(a.keys + wb + q * 32 + lane) ^ (wb + q * 32 + lane);      // the buffer holds one value: make the keys distinct
                else key[q] = nxt[q] ^ (wb + q * 32 + lane);
            }
            unsigned long long k0[ITEMS], v0[ITEMS], k1[ITEMS], v1[ITEMS];
#pragma unroll
            for (int q = 0; q < ITEMS; q++) ldb(bucket(a, p, key[q]), k0[q], v0[q], k1[q], v1[q]);
            if (MODE == 2 && wb + 8 * WT < s0 + SUPER) {
#pragma unroll
                for (int q = 0; q < ITEMS; q++) nxt[q] = __ldcs(a.keys + wb + 8 * WT + q * 32 + lane);
            }
#pragma unroll
            for (int q = 0; q < ITEMS; q++) {
                const unsigned long long r = (k0[q] == key[q]) ? v0[q] : ((k1[q] == key[q]) ? v1[q] : ~0ULL);
                if (MODE == 0) acc += r;
                else __stcs(a.out + wb + q * 32 + lane, r);
            }
        }
    }
    if (MODE == 0 && acc == 0x1234567) a.out[0] = acc;
}

// W: unit u = (partition p, block b), p major. The segment of partition p inside block b starts at b * (seg * n_part) + p * seg.
template <int ITEMS, int MINB>
__global__ void __launch_bounds__(256, MINB) seg_kernel(Args a) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr uint32_t WT = 32 * ITEMS;
    constexpr uint32_t UNITS = 64;                       // units per CTA claim (8 per warp)
    __shared__ unsigned int s_next;
    const uint32_t n_units = a.n_part * a.n_blk;
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) s_next = atomicAdd(a.ctr, 1u);
        __syncthreads();
        const uint32_t u0 = s_next * UNITS;
        if (u0 >= n_units) break;
        for (uint32_t u = u0 + warp; u < u0 + UNITS && u < n_units; u += 8) {
            const uint32_t p = u / a.n_blk, b = u - p * a.n_blk;
            const uint64_t base = ((uint64_t)b * a.n_part + p) * a.seg;
            unsigned long long nxt[ITEMS];
#pragma unroll
            for (int q = 0; q < ITEMS; q++) nxt[q] = (q * 32 + lane < a.seg) ? __ldcs(a.keys + base + q * 32 + lane) : 0ULL;
            for (uint32_t i0 = 0; i0 < a.seg; i0 += WT) {
                unsigned long long key[ITEMS];
#pragma unroll
                for (int q = 0; q < ITEMS; q++) key[q] = nxt[q] ^ (base + i0 + q * 32 + lane);
                unsigned long long k0[ITEMS], v0[ITEMS], k1[ITEMS], v1[ITEMS];
#pragma unroll
                for (int q = 0; q < ITEMS; q++)
                    if (i0 + q * 32 + lane < a.seg) ldb(bucket(a, p, key[q]), k0[q], v0[q], k1[q], v1[q]);
                if (i0 + WT < a.seg) {
#pragma unroll
                    for (int q = 0; q < ITEMS; q++)
                        nxt[q] = (i0 + WT + q * 32 + lane < a.seg) ? __ldcs(a.keys + base + i0 + WT + q * 32 + lane) : 0ULL;
                }
#pragma unroll
                for (int q = 0; q < ITEMS; q++) {
                    if (i0 + q * 32 + lane >= a.seg) continue;
                    const unsigned long long r = (k0[q] == key[q]) ? v0[q] : ((k1[q] == key[q]) ? v1[q] : ~0ULL);
                    __stcs(a.out + base + i0 + q * 32 + lane, r);
                }
            }
        }
    }
}


// T: the streams go through shared memory with bulk-async copies (cp.async.bulk + mbarrier in, cp.async.bulk + bulk groups
// out) in a per-WARP ring of NST stages of CH rows; L1TEX then only sees the table lookups (LDS / STS for keys and results).
__device__ __forceinline__ uint32_t s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
template <int CH, int NST, int WARPS, int MINB>
__global__ void __launch_bounds__(WARPS * 32, MINB) tma_kernel(Args a) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned long long *buf = reinterpret_cast<unsigned long long *>(smem) + (size_t)warp * NST * CH;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + (size_t)WARPS * NST * CH * 8) + warp * NST;
    const uint64_t rows_per_part = a.n / a.n_part;
    const uint32_t n_chunks = (uint32_t)(a.n / CH);
    if (lane == 0) {
        for (int s = 0; s < NST; s++) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(bars + s)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    uint32_t chunk_of[NST];
    uint32_t uses = 0;                                   // chunks consumed so far (stage = uses % NST, parity = (uses / NST) & 1)
    auto issue = [&](int s) -> uint32_t {                // lane 0: claim a chunk and start its load into stage s
        uint32_t c = 0xffffffffu;
        if (lane == 0) {
            c = atomicAdd(a.ctr, 1u);
            if (c < n_chunks) {
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(bars + s)), "r"(CH * 8) : "memory");
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s32(buf + s * CH)),
                             "l"(a.keys + (uint64_t)c * CH), "r"(CH * 8), "r"(s32(bars + s)) : "memory");
            }
        }
        return __shfl_sync(0xffffffffu, c, 0);
    };
#pragma unroll
    for (int s = 0; s < NST - 1; s++) chunk_of[s] = issue(s);
    chunk_of[NST - 1] = 0xffffffffu;
    for (;;) {
        const int s = uses % NST;
        uint32_t c = chunk_of[0];
#pragma unroll
        for (int j = 1; j < NST; j++) if (s == j) c = chunk_of[j];
        if (c >= n_chunks) break;
        {   // wait for the stage's keys
            const uint32_t par = (uses / NST) & 1u;
            uint32_t ok;
            do {
                asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(s32(bars + s)), "r"(par) : "memory");
            } while (!ok);
        }
        unsigned long long *st = buf + s * CH;
        const uint64_t row0 = (uint64_t)c * CH;
        const uint32_t p = (uint32_t)(row0 / rows_per_part) < a.n_part ? (uint32_t)(row0 / rows_per_part) : a.n_part - 1;
#pragma unroll 1
        for (int i0 = 0; i0 < CH; i0 += 128) {
            unsigned long long key[4];
#pragma unroll
            for (int q = 0; q < 2; q++) {
                const ulonglong2 v = *reinterpret_cast<const ulonglong2 *>(st + i0 + (q * 32 + lane) * 2);
                key[2 * q] = v.x ^ (row0 + i0 + (q * 32 + lane) * 2);
                key[2 * q + 1] = v.y ^ (row0 + i0 + (q * 32 + lane) * 2 + 1);
            }
            unsigned long long k0[4], v0[4], k1[4], v1[4];
#pragma unroll
            for (int q = 0; q < 4; q++) ldb(bucket(a, p, key[q]), k0[q], v0[q], k1[q], v1[q]);
            unsigned long long r[4];
#pragma unroll
            for (int q = 0; q < 4; q++) r[q] = (k0[q] == key[q]) ? v0[q] : ((k1[q] == key[q]) ? v1[q] : ~0ULL);
#pragma unroll
            for (int q = 0; q < 2; q++) *reinterpret_cast<ulonglong2 *>(st + i0 + (q * 32 + lane) * 2) = make_ulonglong2(r[2 * q], r[2 * q + 1]);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) {
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(a.out + row0), "r"(s32(st)), "r"(CH * 8) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            // the stage consumed BEFORE this one may be refilled once its store has read it: all groups but the newest
            asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
        }
        __syncwarp();
        uses++;
        const int rs = (uses + NST - 2) % NST;          // the stage consumed before this one ( = the one NOT loaded ahead )
        const uint32_t nc = issue(rs);
#pragma unroll
        for (int j = 0; j < NST; j++) if (rs == j) chunk_of[j] = nc;
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

template <typename F>
static float timeit(F &&f, unsigned int *ctr) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float best = 1e9f;
    for (int rep = 0; rep < 4; rep++) {
        CK(cudaMemset(ctr, 0, 4));
        CK(cudaEventRecord(e0));
        f();
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        CK(cudaGetLastError());
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (rep > 0 && ms < best) best = ms;
    }
    return best;
}

int main() {
    const uint64_t n = 1ull << 29;                      // 512 Mi rows: 4 GB keys, 4 GB out
    unsigned long long *keys, *out; unsigned int *ctr; char *tab;
    CK(cudaMalloc(&keys, n * 8)); CK(cudaMalloc(&out, n * 8)); CK(cudaMalloc(&ctr, 4));
    CK(cudaMemset(keys, 0x5a, n * 8));
    const uint32_t nb_part = 930000;                    // ~30 MB per partition (781 K keys at load factor 0.42)
    const uint32_t P = 128;
    CK(cudaMalloc(&tab, (size_t)P * nb_part * 32));
    CK(cudaMemset(tab, 0, (size_t)P * nb_part * 32));
    cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
    const int sms = prop.multiProcessorCount;
    printf("device %s, %d SMs, %llu rows per launch\n", prop.name, sms, (unsigned long long)n);
    auto report = [&](const char *name, float ms) {
        printf("%-44s %8.3f ms  %7.1f G rows/s  %5.2f clk/row/SM\n", name, ms, n / ms / 1e6, ms * 1e-3 * 1.965e9 * sms / n);
        fflush(stdout);
    };
    for (int regime = 0; regime < 2; regime++) {
        Args a{};
        a.tab = tab; a.keys = keys; a.out = out; a.n = n; a.ctr = ctr;
        a.nb = regime == 0 ? (1u << 20) : nb_part;
        a.n_part = regime == 0 ? 1 : P;
        const char *rg = regime == 0 ? "L2 " : "PT ";
        char nm[128];
#define RUN_STREAM(ITEMS, MODE, MINB, label)                                                                        \
        { snprintf(nm, sizeof nm, "%s%s items=%d ctas/sm=%d", rg, label, ITEMS, MINB);                              \
          report(nm, timeit([&] { stream_kernel<ITEMS, MODE, MINB><<<sms * MINB, 256>>>(a); }, ctr)); }
        RUN_STREAM(8, 0, 4, "G gather only      ")
        RUN_STREAM(4, 1, 4, "S stream           ")
        RUN_STREAM(8, 1, 4, "S stream           ")
        RUN_STREAM(4, 2, 4, "P stream+prefetch  ")
        RUN_STREAM(8, 2, 3, "P stream+prefetch  ")

#define RUN_TMA(CH, NST, WARPS, MINB)                                                                                 \
        { const size_t sm = (size_t)WARPS * NST * CH * 8 + WARPS * NST * 8;                                                \
          CK(cudaFuncSetAttribute(tma_kernel<CH, NST, WARPS, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm)); \
          snprintf(nm, sizeof nm, "%sT bulk-async ch=%d st=%d warps=%d ctas/sm=%d", rg, CH, NST, WARPS, MINB);          \
          report(nm, timeit([&] { tma_kernel<CH, NST, WARPS, MINB><<<sms * MINB, WARPS * 32, sm>>>(a); }, ctr)); }
        RUN_TMA(512, 3, 8, 2)
        RUN_TMA(512, 3, 4, 4)
        RUN_TMA(256, 3, 8, 4)
        RUN_TMA(256, 3, 4, 8)
        RUN_TMA(256, 4, 4, 6)
        RUN_TMA(128, 4, 8, 6)
        RUN_TMA(1024, 2, 4, 3)
        for (uint32_t seg : {128u, 512u}) {
            a.seg = seg;
            a.n_blk = (uint32_t)(n / ((uint64_t)seg * a.n_part));
            if (regime == 0) { a.n_part = P; a.n_blk = (uint32_t)(n / ((uint64_t)seg * P)); }   // same walk, one shared 32 MB table
            Args w = a;
            if (regime == 0) w.nb = (1u << 20) / P;      // P tables of 256 KB: the whole set is L2-resident
            snprintf(nm, sizeof nm, "%sW segments of %u rows items=4 ctas/sm=4", rg, seg);
            report(nm, timeit([&] { seg_kernel<4, 4><<<sms * 4, 256>>>(w); }, ctr));
            if (regime == 0) a.n_part = 1;
        }
    }
    // plain copy of the same volume for reference (8 B in + 8 B out per row)
    {
        cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
        float best = 1e9f;
        for (int r = 0; r < 4; r++) {
            CK(cudaEventRecord(e0));
            CK(cudaMemcpyAsync(out, keys, n * 8, cudaMemcpyDeviceToDevice));
            CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
            float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best) best = ms;
        }
        printf("%-44s %8.3f ms  %7.1f G rows/s  (%.0f GB/s)\n", "cudaMemcpy D2D 8 B in + 8 B out", best, n / best / 1e6, n * 16 / best / 1e6);
    }
    return 0;
}
