#!/bin/bash
# parity tests (all, not -x), bench with both store modes, ncu launch list + full capture of the fused level kernel
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_t.json 2> gpurun_out/bench_t.err; echo "bench exit $?"; cat gpurun_out/bench_t.json
timeout 600 python bench.py --no-cpu-baseline --row-store > gpurun_out/bench_r.json 2> gpurun_out/bench_r.err; echo "bench exit $?"; cat gpurun_out/bench_r.json
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
timeout 600 $B > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_update_level|k_sample_tracks" -s 18 -c 9 -o gpurun_out/prof_b $B > gpurun_out/ncu_b.log 2>&1
ls -la gpurun_out
