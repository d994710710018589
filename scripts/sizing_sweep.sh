#!/bin/bash
# SURVEY §8d "sizing sweep": 1 B uniform probe rows against unique int64 build sides of 2^16 .. 2^27 keys (table 2 MB .. 4 GB
# at load factor 0.5, crossing the 126 MB L2), un-partitioned (direct) vs radix-partitioned vs what layout=auto picks.
FMT='import json,sys; d=json.loads(sys.stdin.read()); print({k:(round(v,2) if isinstance(v,float) else v) for k,v in d.items() if k in ("layout","radix_bits","table_MB","build_ms","probe_part_ms","probe_kernel_ms","probe_ms","probe_Gtps","probe_roofline_frac_36B")})'
for lg in 16 18 20 21 22 23 24 25 26 27; do
  for l in direct radix auto; do
    echo -n "build=2^$lg layout=$l "; python scripts/quick_bench.py --build $((1 << lg)) --layout $l --reps 2 --chunk 1000000000 2>&1 | tail -1 | python -c "$FMT"
  done
done
