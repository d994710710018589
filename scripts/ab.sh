#!/bin/bash
# usage: scripts/ab.sh <rounds> <lib or -> ...   C2 shape (100 M x 1 B), library defaults (load factor, layout auto), per-phase ms
FMT='import json,sys; d=json.loads(sys.stdin.read()); print({k:round(d[k],3) for k in ("build_copy_ms","build_part_ms","build_insert_ms","probe_part_ms","probe_kernel_ms")}, d["matches"])'
rounds=$1; shift
for r in $(seq $rounds); do
  for lib in "$@"; do
    if [ "$lib" = "-" ]; then unset B200J_LIB; else export B200J_LIB=$PWD/gpurun_variants/$lib.so; fi
    echo -n "$lib "; python scripts/quick_bench.py --layout auto --lf 0 --chunk 1000000000 --reps 3 ${AB_ARGS} | tail -1 | python -c "$FMT"
  done
done
