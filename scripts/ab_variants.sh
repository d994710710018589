#!/bin/bash
# usage: scripts/ab_variants.sh <rounds> <lib or -> ...   ("-" = the in-tree library); C2 shape, radix layout, per-phase ms
FMT='import json,sys; d=json.loads(sys.stdin.read()); print({k:round(d[k],3) for k in ("build_part_ms","build_insert_ms","probe_part_ms","probe_kernel_ms")})'
rounds=$1; shift
for r in $(seq $rounds); do
  for lib in "$@"; do
    if [ "$lib" = "-" ]; then unset B200J_LIB; else export B200J_LIB=$PWD/$lib; fi
    echo -n "$lib "; python scripts/quick_bench.py --layout radix --chunk 1000000000 --reps 3 | tail -1 | python -c "$FMT"
  done
done
