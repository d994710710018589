#!/bin/bash
# round-2 session M, GPU call 1: full GPU suite with the partition-order gather, ILP insert suite, A/B bench lines, launch list
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/m1_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/m1_pytest.log
tail -5 gpurun_out/m1_pytest.log
B200J_INS_ILP=1 python -m pytest tests/test_gpu_join.py tests/test_gpu_dropin.py tests/test_tpch.py -m gpu -x -q > gpurun_out/m1_pytest_ilp.log 2>&1; echo "pytest rc $?" >> gpurun_out/m1_pytest_ilp.log
tail -3 gpurun_out/m1_pytest_ilp.log
python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e > gpurun_out/m1_bench_default.json 2> gpurun_out/m1_bench_default.err
B200J_GATHER_SEG=0 python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/m1_bench_noseg.json 2> gpurun_out/m1_bench_noseg.err
B200J_INS_ILP=1 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e --no-gather > gpurun_out/m1_bench_ilp.json 2> gpurun_out/m1_bench_ilp.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/m1_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e --no-verify > gpurun_out/m1_ncu.log 2>&1
python - <<'PY'
import json
for f in ("default","noseg","ilp"):
    try:
        d=json.loads(open(f"gpurun_out/m1_bench_{f}.json").read().strip().splitlines()[-1])
        print(f, d["ms_per_step"], d["phases"]["ms_per_step"], d["phases"].get("gather"), d.get("verified"))
    except Exception as e:
        print(f, "ERR", e)
PY
