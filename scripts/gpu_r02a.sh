#!/bin/bash
# round 2, first call: parity tests, A/B of the one-launch top of the hierarchy and of the register cap, launch list + full captures
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
export VARIANTS="${VARIANTS:-;HGB_TOP_BUDGET=0|;HGB_TOP_BUDGET=250000|;HGB_LEVEL_MINB=5|;--workload c2;HGB_TOP_BUDGET=0|--workload c2;--workload c5}"
bash scripts/gpu_ab.sh
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_update_level|k_update_top" -s 6 -c 3 -o gpurun_out/prof_update_level $B > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out
