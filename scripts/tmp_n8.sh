#!/bin/bash
mkdir -p gpurun_out
bash scripts/n_bench.sh 8 2>&1 | tee gpurun_out/m17_n8.txt; cp gpurun_out/nb.out gpurun_out/m17_n8_full.out
bash scripts/n_bench.sh 4 2>&1 | tee gpurun_out/m17_n4.txt; cp gpurun_out/nb.out gpurun_out/m17_n4_full.out
