#!/bin/bash
mkdir -p gpurun_out
for lib in - slotexp1 slotexp2; do
  if [ "$lib" = "-" ]; then unset B200J_LIB; else export B200J_LIB=$PWD/gpurun_variants/$lib.so; fi
  echo -n "$lib "; python scripts/bench_hint_probe.py | tail -1
done 2>&1 | tee gpurun_out/m5_slot_exp.txt
