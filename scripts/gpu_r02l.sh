#!/bin/bash
# A/B at N ranks: persistent CTAs of the tail's compaction. usage: N=2 bash scripts/gpu_r02l.sh
mkdir -p gpurun_out
N=${N:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
for m in ${MODES:-0 148}; do
  HGB_COMPACT_CTAS=$m timeout 600 $TR bench.py --gpus $N --no-cpu-baseline --no-facade --steps 200 --warmup 5 > gpurun_out/ab_compact_${m}_n$N.json 2> gpurun_out/ab_compact_${m}_n$N.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/ab_compact_${m}_n$N.json") if l.startswith("{")][-1])
    print("N=$N ctas $m: ms/step %.4f e2e %.0f exch %s"%(d["ms_per_step"],d["e2e"]["value"],d["exchange_check"]["ok"]), {k:round(v,4) for k,v in d["roofline"]["kernel_ms_per_step"].items()})
except Exception as e:
    print("failed", e); print(open("gpurun_out/ab_compact_${m}_n$N.err").read()[-1200:])
PY
done
