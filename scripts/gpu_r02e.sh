#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
run() { n=$1; shift; timeout 900 env $ENVV python bench.py "$@" > gpurun_out/$n.json 2> gpurun_out/$n.err; echo "$n exit $?"; head -c 2500 gpurun_out/$n.json; echo; tail -3 gpurun_out/$n.err; }
run bench_c4
ENVV="HGB_L2_FETCH=64" run bench_c4_l2f64 --no-cpu-baseline --no-facade
ENVV="HGB_L2_FETCH=32" run bench_c4_l2f32 --no-cpu-baseline --no-facade
ENVV="HGB_LEVEL_BLOCK=128" run bench_c4_b128 --no-cpu-baseline --no-facade
run bench_c1 --workload c1
run bench_c2 --workload c2
run bench_c3 --workload c3
run bench_c5 --workload c5
timeout 900 python bench.py --impl reference --steps 6 --warmup 2 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; cat gpurun_out/bench_ref.json
ls -la gpurun_out | head -4
