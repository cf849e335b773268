#!/bin/bash
mkdir -p gpurun_out
export VARIANTS="${VARIANTS:-;HGB_ASYNC_POST=0|;HGB_LEVEL_BLOCK=128|;HGB_LEVEL_BLOCK=128 HGB_ASYNC_POST=0|;--lookahead 113664;--lookahead -1;--workload c2;--workload c5}"
bash scripts/gpu_ab.sh
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_update_level|k_update_top" -s 6 -c 3 -o gpurun_out/prof_update_level $B > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out | head -5
