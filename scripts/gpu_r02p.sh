#!/bin/bash
# c4 weak at N ranks with the frame mode chosen per frame (default): value from back-to-back updates, e2e with a read after every update
mkdir -p gpurun_out
N=${N:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus $N --no-cpu-baseline --steps 100 --warmup 5 > gpurun_out/bench_c4_n$N.json 2> gpurun_out/bench_c4_n$N.err; echo "exit $?"
python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/bench_c4_n$N.json") if l.startswith("{")][-1])
    print("N=$N ms/step %.4f value %.0f e2e %.0f facade %s exch %s"%(d["ms_per_step"],d["value"],d["e2e"]["value"],d.get("e2e_facade",{}).get("value"),d["exchange_check"]["ok"]), {k:round(v,4) for k,v in d["roofline"]["kernel_ms_per_step"].items()})
except Exception as e:
    print("failed", e); print(open("gpurun_out/bench_c4_n$N.err").read()[-1500:])
PY
