#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -v "^  File" gpurun_out/pytest_gpu.log | tail -6
timeout 600 python scripts/edit_cost.py > gpurun_out/edit_cost.json 2> gpurun_out/edit_cost.err; cat gpurun_out/edit_cost.json; tail -2 gpurun_out/edit_cost.err
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench exit $?"; python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/bench_c4.json") if l.startswith("{")][-1])
print("ms/step %.4f value %.0f e2e %.0f facade %s frac %.3f"%(d["ms_per_step"],d["value"],d["e2e"]["value"],d.get("e2e_facade",{}).get("value"),d["roofline"]["frac"]), d["roofline"]["kernel_ms_per_step"])
PY
