#!/bin/bash
# usage: scripts/sweep.sh "<lib or ->" "<bits list>" [extra quick_bench args]
FMT='import json,sys; d=json.loads(sys.stdin.read()); print({k:(round(v,2) if isinstance(v,float) else v) for k,v in d.items() if k in ("radix_bits","build_part_ms","build_insert_ms","probe_part_ms","probe_kernel_ms","probe_wall_ms","build_wall_ms","table_MB")})'
lib=$1; bits=$2; shift 2
for b in $bits; do
  if [ "$lib" != "-" ]; then export B200J_LIB=$lib; fi
  echo -n "$lib bits=$b "; python scripts/quick_bench.py --layout radix --radix-bits $b --reps 2 --chunk 1000000000 "$@" 2>&1 | tail -1 | python -c "$FMT"
done
