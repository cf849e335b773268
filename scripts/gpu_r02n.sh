#!/bin/bash
# A/B: visible list streamed to pinned host memory by the compaction kernel (HGB_HOST_STREAM), exact-NV multi-view compaction
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for m in 0 1 0 1; do
  HGB_HOST_STREAM=$m timeout 600 python bench.py --no-cpu-baseline --no-facade --steps 100 > gpurun_out/ab_stream_$m.json 2> gpurun_out/ab_stream_$m.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/ab_stream_$m.json") if l.startswith("{")][-1])
    print("stream $m: ms/step %.4f e2e %.0f (%.4f ms) full_upload %.0f"%(d["ms_per_step"],d["e2e"]["value"],d["e2e"]["ms_per_step"],d["e2e_full_upload"]["value"]), {k:round(v,4) for k,v in d["roofline"]["kernel_ms_per_step"].items()})
except Exception as e:
    print("failed", e); print(open("gpurun_out/ab_stream_$m.err").read()[-1500:])
PY
done
for w in c2 c5; do
  timeout 600 python bench.py --workload $w --no-cpu-baseline --no-facade > gpurun_out/ab_stream_$w.json 2> gpurun_out/ab_stream_$w.err; python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/ab_stream_$w.json") if l.startswith("{")][-1]); print("$w ms/step %.4f e2e %.0f"%(d["ms_per_step"],d["e2e"]["value"]))
PY
done
timeout 600 python scripts/multiview_compact.py > gpurun_out/multiview_onepass.json 2> gpurun_out/multiview.err; cat gpurun_out/multiview_onepass.json; tail -2 gpurun_out/multiview.err
