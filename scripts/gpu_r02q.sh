#!/bin/bash
# A/B: registers of the level kernel (HGB_LEVEL_MINB: 6 = 40 registers / 48 warps per SM, 8 = 32 registers / 60 warps, 5 = 48 registers)
mkdir -p gpurun_out
HGB_LEVEL_MINB=8 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "c1_static or position_tracks or multi_view or c3_small" 2>&1 | tail -2
for m in 6 8 6 8; do
  HGB_LEVEL_MINB=$m timeout 600 python bench.py --no-cpu-baseline --no-facade --steps 100 > gpurun_out/ab_minb_$m.json 2> gpurun_out/ab_minb_$m.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/ab_minb_$m.json") if l.startswith("{")][-1])
    print("minb $m: ms/step %.4f e2e %.0f frac %.3f"%(d["ms_per_step"],d["e2e"]["value"],d["roofline"]["frac"]), {k:round(v,4) for k,v in d["roofline"]["kernel_ms_per_step"].items()})
except Exception as e:
    print("failed", e); print(open("gpurun_out/ab_minb_$m.err").read()[-1500:])
PY
done
