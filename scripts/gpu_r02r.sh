#!/bin/bash
# A/B: L2 look-ahead distance of the level kernel (positions; default = 148 x 6 x 256 = 227328)
mkdir -p gpurun_out
for m in 0 113664 454656 909312; do
  timeout 600 python bench.py --no-cpu-baseline --no-facade --steps 100 --lookahead $m > gpurun_out/ab_la_$m.json 2> gpurun_out/ab_la_$m.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/ab_la_$m.json") if l.startswith("{")][-1])
    print("lookahead $m: ms/step %.4f frac %.3f"%(d["ms_per_step"],d["roofline"]["frac"]), {k:round(v,4) for k,v in d["roofline"]["kernel_ms_per_step"].items()})
except Exception as e:
    print("failed", e); print(open("gpurun_out/ab_la_$m.err").read()[-1500:])
PY
done
