#!/bin/bash
set -x
mkdir -p gpurun_out
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/m4_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e --no-verify > gpurun_out/m4_ncu.log 2>&1
grep -c . gpurun_out/m4_launches.csv
