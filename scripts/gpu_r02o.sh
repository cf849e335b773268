#!/bin/bash
# A/B: asynchronous tail chosen per frame (default) vs forced on (HGB_ASYNC_POST=1) vs off (0): e2e legs
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for m in 1 2 0 1 2; do
  HGB_ASYNC_POST=$m timeout 600 python bench.py --no-cpu-baseline --steps 100 > gpurun_out/ab_auto_$m.json 2> gpurun_out/ab_auto_$m.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/ab_auto_$m.json") if l.startswith("{")][-1])
    print("async $m: ms/step %.4f e2e %.0f (%.4f ms) facade %.0f"%(d["ms_per_step"],d["e2e"]["value"],d["e2e"]["ms_per_step"],d["e2e_facade"]["value"]))
except Exception as e:
    print("failed", e); print(open("gpurun_out/ab_auto_$m.err").read()[-1500:])
PY
done
