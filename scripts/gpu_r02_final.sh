#!/bin/bash
# Round-2 evidence, one GPU: parity tests, smoke, both bench arms, every workload, then — only after the same bench command exited 0
# without ncu — the ncu launch list and the full captures of the hot kernels.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
nproc >> gpurun_out/gpu.txt; free -g >> gpurun_out/gpu.txt; lscpu | egrep "Model name|^CPU\(s\)|Thread|Socket" >> gpurun_out/gpu.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -v "^  File" gpurun_out/pytest_gpu.log | tail -5
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke.log
timeout 700 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; cat gpurun_out/bench.json; tail -3 gpurun_out/bench.err
timeout 600 python bench.py --impl reference --steps 6 --warmup 2 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; cat gpurun_out/bench_ref.json
for w in c1 c2 c3 c5; do
  timeout 600 python bench.py --workload $w > gpurun_out/bench_${w}_n1.json 2> gpurun_out/bench_${w}_n1.err; echo "$w exit $?"
done
timeout 600 python scripts/edit_cost.py > gpurun_out/edit_cost.json 2> gpurun_out/edit_cost.err
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
timeout 600 $B > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_update_level|k_update_top" -s 18 -c 20 -o gpurun_out/prof_update_level $B > gpurun_out/ncu2.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_compact|k_sample_tracks" -s 6 -c 4 -o gpurun_out/prof_others $B > gpurun_out/ncu3.log 2>&1
ls -la gpurun_out
