#!/bin/bash
mkdir -p gpurun_out
AB_ARGS="" bash scripts/ab.sh 2 - inswarp 2>&1 | tee gpurun_out/m7_inswarp.txt
B200J_LIB=$PWD/gpurun_variants/inswarp.so python -m pytest tests/test_gpu_join.py -m gpu -x -q 2>&1 | tail -3 | tee -a gpurun_out/m7_inswarp.txt
