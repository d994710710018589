#!/bin/bash
# End-of-round validation on one B200 (run under gpurun): GPU suite, smoke, the contract bench line + reference arm,
# ncu launch list and one ncu --set full capture of the library's kernels (summarised with scripts/ncu_summary.py).
# usage: scripts/final_validation.sh <tag> [ncu-only | no-ncu-full]     -> gpurun_out/<tag>_*
tag=${1:-final}
mkdir -p gpurun_out
if [ "$2" != "ncu-only" ]; then
python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/${tag}_pytest.log
tail -3 gpurun_out/${tag}_pytest.log
python __graft_entry__.py smoke > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke rc $?" >> gpurun_out/${tag}_smoke.log
tail -2 gpurun_out/${tag}_smoke.log
python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc $?"
python bench.py --impl reference > gpurun_out/${tag}_bench_reference_arm.json 2> gpurun_out/${tag}_bench_reference_arm.err; echo "reference arm rc $?"
fi
ncu --metrics gpu__time_duration.sum --clock-control none -k 'regex:^(build_insert|csr_|gather_|part_|probe_|slot_order|compact_pairs|narrow_)' -c 200 --csv --log-file gpurun_out/${tag}_launches_bench_C2.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-verify > gpurun_out/${tag}_ncu_launches.log 2>&1; echo "launch list rc $?"
if [ "$2" != "no-ncu-full" ]; then
# the .ncu-rep stays on the box (hundreds of MB with --import-source): only its per-kernel summary comes back
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:^(build_insert|gather_|part_block|part_scatter|probe_|slot_order)' -c 70 -f -o /tmp/${tag}_full \
    python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e --no-verify > gpurun_out/${tag}_ncu_full.log 2>&1; echo "full capture rc $?"
python scripts/ncu_summary.py /tmp/${tag}_full.ncu-rep > gpurun_out/${tag}_ncu_full_bench_C2.txt 2>> gpurun_out/${tag}_ncu_full.log
fi
ls -la gpurun_out/${tag}_* | awk '{print $5, $9}'
[ "$2" = "ncu-only" ] && exit 0
python - <<PY
import json
d=json.loads(open("gpurun_out/${tag}_bench.json").read().strip().splitlines()[-1])
print("ms/step", d["ms_per_step"], "value", d["value"], "verified", d["verified"], "e2e", d["e2e"] and d["e2e"]["ms_per_step"], "roofline", d["roofline"]["frac"], d["roofline"]["traffic"])
print("gather", {k: d["phases"]["gather"][k] for k in ("ms", "mode", "gather_roofline_frac", "frac_on_survey_60B", "hint_extra_probe_ms", "verified")})
print("cpu", d["cpu_baseline"] and {k: d["cpu_baseline"][k] for k in ("value", "cores", "kind")})
PY
