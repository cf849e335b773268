#!/bin/bash
# parity tests, then the secondary-kernel roofline probe
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 900 python scripts/kernel_rooflines.py > gpurun_out/kernel_rooflines.json 2> gpurun_out/kernel_rooflines.err; echo "probe exit $?"
cat gpurun_out/kernel_rooflines.json; tail -5 gpurun_out/kernel_rooflines.err
