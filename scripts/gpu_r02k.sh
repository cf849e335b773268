#!/bin/bash
# A/B: persistent CTAs of the tail's compaction
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "append or async_tail or multi_view or c1_static" 2>&1 | tail -3
for m in 0 148 296 74; do
  HGB_COMPACT_CTAS=$m timeout 600 python bench.py --no-cpu-baseline --no-facade --steps 200 > gpurun_out/ab_compact_$m.json 2> gpurun_out/ab_compact_$m.err
  python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/ab_compact_$m.json") if l.startswith("{")][-1])
print("ctas $m: ms/step %.4f e2e %.0f"%(d["ms_per_step"],d["e2e"]["value"]), {k:round(v,4) for k,v in d["roofline"]["kernel_ms_per_step"].items()})
PY
done
