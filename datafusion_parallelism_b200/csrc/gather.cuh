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
