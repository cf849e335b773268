#!/bin/bash
mkdir -p gpurun_out
export VARIANTS="${VARIANTS:-;HGB_LEVEL_BLOCK=128|;HGB_TOP_BUDGET=250000|;--workload c2;--workload c5}"
bash scripts/gpu_ab.sh
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_update_level|k_update_top|k_sample_tracks" -s 8 -c 4 -o gpurun_out/prof_update_level $B > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out | head -5
