#!/bin/bash
# A/B at N ranks: frame mode forced async (1) vs chosen per frame (2) on c5 strong and c4 (e2e leg)
mkdir -p gpurun_out
N=${N:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
for w in "--workload c5 --strong" ${ALSO_C4:+""}; do for m in 1 2; do
  HGB_ASYNC_POST=$m timeout 500 $TR bench.py --gpus $N --no-cpu-baseline --no-facade --steps 100 --warmup 5 $w > gpurun_out/ab_mode.json 2> gpurun_out/ab_mode.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/ab_mode.json") if l.startswith("{")][-1])
    print("N=$N [$w] async=$m: ms/step %.4f e2e %.0f (%.4f ms)"%(d["ms_per_step"],d["e2e"]["value"],d["e2e"]["ms_per_step"]))
except Exception as e:
    print("failed", e); print(open("gpurun_out/ab_mode.err").read()[-1500:])
PY
done; done
