#!/bin/bash
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/m2_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/m2_pytest.log
tail -4 gpurun_out/m2_pytest.log
bash scripts/ab_variants.sh 2 gpurun_variants/base.so gpurun_variants/pf1.so gpurun_variants/pf2.so gpurun_variants/ipf1.so gpurun_variants/ipf1pf1.so 2>&1 | tee gpurun_out/m2_ab.txt
