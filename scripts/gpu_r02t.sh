#!/bin/bash
# one bench line at N ranks: N=8 ARGS="--workload c5 --strong" NAME=bench_c5_strong bash scripts/gpu_r02t.sh
mkdir -p gpurun_out
N=${N:-8}; NAME=${NAME:-bench_c5_strong}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 500 $TR bench.py --gpus $N --no-cpu-baseline --steps 100 --warmup 5 $ARGS > gpurun_out/${NAME}_n$N.json 2> gpurun_out/${NAME}_n$N.err; echo "exit $?"
python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/${NAME}_n$N.json") if l.startswith("{")][-1])
    print("N=$N ms/step %.4f value %.0f e2e %.0f exch %s frac %.3f"%(d["ms_per_step"],d["value"],d["e2e"]["value"],d["exchange_check"]["ok"],d["roofline"]["frac"]), {k:round(v,4) for k,v in d["roofline"]["kernel_ms_per_step"].items()})
except Exception as e:
    print("failed", e); print(open("gpurun_out/${NAME}_n$N.err").read()[-1500:])
PY
