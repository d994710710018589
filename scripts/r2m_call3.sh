#!/bin/bash
set -x
mkdir -p gpurun_out
python -m pytest tests/test_gather.py -m gpu -x -q > gpurun_out/m3_pytest_gather.log 2>&1; echo "pytest rc $?" >> gpurun_out/m3_pytest_gather.log
tail -4 gpurun_out/m3_pytest_gather.log
python -m pytest tests -m gpu -x -q > gpurun_out/m3_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/m3_pytest.log
tail -4 gpurun_out/m3_pytest.log
python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e > gpurun_out/m3_bench.json 2> gpurun_out/m3_bench.err
tail -3 gpurun_out/m3_bench.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/m3_bench.json").read().strip().splitlines()[-1])
print(d["ms_per_step"], d["phases"]["ms_per_step"], d.get("verified"))
print(json.dumps(d["phases"].get("gather"), indent=1))
PY
ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/m3_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e --no-verify > gpurun_out/m3_ncu.log 2>&1
grep -c . gpurun_out/m3_launches.csv
