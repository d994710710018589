// radix.cuh — shared-memory-staged radix partitioning (histogram + scatter) in front of build/probe.
#pragma once
#include "kernels.cuh"
namespace b200j {
struct RadixScratch {
    int dummy = 0;
};
}  // namespace b200j
