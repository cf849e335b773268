#!/bin/bash
# A/B: budget of the one-launch top of the hierarchy on the workloads it was not tuned on (c5, c3, c5 at the N=8 strong share)
mkdir -p gpurun_out
for w in c5 c3; do for b in 250000 500000 250000 500000; do
  HGB_TOP_BUDGET=$b timeout 600 python bench.py --workload $w --no-cpu-baseline --no-facade --steps 200 > gpurun_out/ab_top_${w}_$b.json 2> gpurun_out/ab_top_${w}_$b.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/ab_top_${w}_$b.json") if l.startswith("{")][-1])
    print("$w budget $b: ms/step %.4f frac %.3f launches %.0f"%(d["ms_per_step"],d["roofline"]["frac"],d["roofline"]["launches_per_step"]), {k:round(v,4) for k,v in d["roofline"]["kernel_ms_per_step"].items()})
except Exception as e:
    print("failed", e); print(open("gpurun_out/ab_top_${w}_$b.err").read()[-1500:])
PY
done; done
