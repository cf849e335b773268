#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -v "^  File" gpurun_out/pytest_gpu.log | tail -6
export VARIANTS="${VARIANTS:-;HGB_TOP_BUDGET=0|;HGB_TOP_BUDGET=250000|;--workload c2;--workload c3;--workload c5}"
sed -i 's/--no-cpu-baseline \$v/--no-cpu-baseline --no-facade $v/' scripts/gpu_ab.sh
bash scripts/gpu_ab.sh 2>&1 | grep -v "passed\|pytest"
