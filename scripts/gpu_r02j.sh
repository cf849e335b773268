#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -v "^  File" gpurun_out/pytest_gpu.log | tail -25
timeout 600 python scripts/edit_cost.py > gpurun_out/edit_cost.json 2> gpurun_out/edit_cost.err; cat gpurun_out/edit_cost.json; tail -2 gpurun_out/edit_cost.err
