// Micro-benchmark: random 32 B sector reads (one hash-table bucket per lookup) from a table of a given size,
// with a streaming 8 B key read and 12 B pair write beside it — the memory shape of the direct-layout probe.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)
__device__ __forceinline__ uint64_t mix(uint64_t x) { x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33; return x; }
template <int ITEMS, bool STREAM>
__global__ void __launch_bounds__(256) gather(const ulonglong4 *__restrict__ tab, uint64_t mask, const uint64_t *__restrict__ keys, uint64_t n,
                                              uint64_t *__restrict__ outb, uint32_t *__restrict__ outp) {
    for (uint64_t base = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * ITEMS; base < n; base += (uint64_t)gridDim.x * blockDim.x * ITEMS) {
        uint64_t k[ITEMS];
#pragma unroll
        for (int q = 0; q < ITEMS; q += 2) {
            if (STREAM) { ulonglong2 v = __ldcs((const ulonglong2 *)(keys + base + q)); k[q] = v.x; k[q + 1] = v.y; }
            else { k[q] = mix(base + q); k[q + 1] = mix(base + q + 1); }
        }
        unsigned long long a[ITEMS], b[ITEMS], c[ITEMS], d[ITEMS];
#pragma unroll
        for (int q = 0; q < ITEMS; q++) {
            const ulonglong4 *p = tab + (mix(k[q]) & mask);
            asm volatile("ld.global.L1::no_allocate.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(a[q]), "=l"(b[q]), "=l"(c[q]), "=l"(d[q]) : "l"(p));
        }
#pragma unroll
        for (int q = 0; q < ITEMS; q++) {
            uint64_t r = a[q] ^ b[q] ^ c[q] ^ d[q];
            if (STREAM) { __stcs((unsigned long long *)outb + base + q, r); __stcs(outp + base + q, (uint32_t)(base + q)); }
            else if (r == 0x1234567) outb[0] = r;
        }
    }
}
int main(int argc, char **argv) {
    uint64_t n = 1ull << 28;
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    uint64_t *keys, *outb; uint32_t *outp;
    CK(cudaMalloc(&keys, n * 8)); CK(cudaMalloc(&outb, n * 8)); CK(cudaMalloc(&outp, n * 4));
    CK(cudaMemset(keys, 1, n * 8));
    for (int lg = 19; lg <= 27; lg += 2) {        // buckets: 2^lg * 32 B  (16 MB .. 4 GB)
        uint64_t nb = 1ull << lg; ulonglong4 *tab; CK(cudaMalloc(&tab, nb * 32)); CK(cudaMemset(tab, 0, nb * 32));
        for (int stream = 0; stream < 2; stream++) {
            float best = 1e9;
            for (int rep = 0; rep < 4; rep++) {
                CK(cudaEventRecord(e0));
                if (stream) gather<4, true><<<148 * 8, 256>>>(tab, nb - 1, keys, n, outb, outp);
                else gather<4, false><<<148 * 8, 256>>>(tab, nb - 1, keys, n, outb, outp);
                CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
                float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best) best = ms;
            }
            printf("table %7.1f MB  %s  %.3f ms  %.1f Glookups/s  sector-read %.0f GB/s  total(32+%d B) %.0f GB/s\n", nb * 32 / 1e6, stream ? "stream+gather" : "gather-only  ",
                   best, n / best / 1e6, n * 32 / best / 1e6, stream ? 20 : 0, n * (32 + (stream ? 20 : 0)) / best / 1e6);
        }
        CK(cudaFree(tab));
    }
    return 0;
}
