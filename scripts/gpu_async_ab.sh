#!/bin/bash
# N-GPU A/B of the asynchronous frame tail (hgb_options.async_post) with the peer gather, workload c4
N=${N:-8}
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521"
HGB_ASYNC_POST=1 timeout 600 $T tests/mgpu_peer_check.py > gpurun_out/peer_check_async_n$N.log 2>&1; echo "peer check (async tail) N=$N exit $?"; grep "identical" gpurun_out/peer_check_async_n$N.log
for A in 1 0; do
  timeout 900 $T bench.py --gpus $N --steps 30 --warmup 5 --async-post $A --no-cpu-baseline > gpurun_out/bench_c4_async${A}_n$N.json 2> gpurun_out/bench_c4_async${A}_n$N.err
  echo "c4 async=$A N=$N exit $?"
  python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/bench_c4_async${A}_n$N.json") if l.startswith("{")][-1])
    print("  ms/step %.4f value %.0f e2e %.0f resident %.0f launches %d"%(d["ms_per_step"],d["value"],d["e2e"]["value"],d["e2e_resident"]["value"],d["gpu_launches"]), d["roofline"]["kernel_ms_per_step"])
except Exception as e:
    print("  failed", e); print(open("gpurun_out/bench_c4_async${A}_n$N.err").read()[-1500:])
PY
done
