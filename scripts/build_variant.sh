#!/bin/bash
# usage: scripts/build_variant.sh <name> [-D...]   ->  gpurun_variants/<name>.so  (select it with B200J_LIB=$PWD/gpurun_variants/<name>.so)
name=$1; shift
mkdir -p gpurun_variants
exec /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared -cudart static \
  -diag-suppress 177 "$@" -o gpurun_variants/$name.so datafusion_parallelism_b200/csrc/b200join.cu
