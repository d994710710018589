FMT='import json,sys; d=json.loads(sys.stdin.read()); print({k:(round(v,2) if isinstance(v,float) else v) for k,v in d.items() if k in ("layout","radix_bits","table_MB","build_ms","probe_part_ms","probe_kernel_ms","probe_ms","probe_Gtps")})'
for n in 65536 1048576 2097152 2500000 3000000 3500000 4194304 6000000 8388608; do
  for l in direct radix; do
    echo -n "build=$n layout=$l "; python scripts/quick_bench.py --build $n --layout $l --reps 2 --chunk 1000000000 2>&1 | tail -1 | python -c "$FMT"
  done
done
