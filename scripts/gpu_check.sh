#!/bin/bash
# parity tests (all), bench, then — only after the same bench command exited 0 — the ncu launch list and a full capture
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 600 python bench.py ${BENCH_ARGS} > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
if [ -n "$NCU" ]; then
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
timeout 600 $B > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_update_level|k_sample_tracks|k_compact" -s 20 -c 10 -o gpurun_out/prof_full $B > gpurun_out/ncu2.log 2>&1
fi
ls -la gpurun_out | head -30
