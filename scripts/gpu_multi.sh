#!/bin/bash
# N-GPU run: parity of the peer-memory gather against one context (tests/mgpu_peer_check.py), then the bench (own arm) through torchrun
# with both exchanges: peer (default) and nccl (A/B), workloads c4 (one camera per rank batch) and c5 (one camera per instance)
N=${N:-2}
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $T tests/mgpu_peer_check.py > gpurun_out/peer_check_n$N.log 2>&1; echo "peer check N=$N exit $?"; grep -v "^\*\|OMP_NUM\|^$" gpurun_out/peer_check_n$N.log | tail -12
for W in ${WORKLOADS:-c4 c5}; do
  for X in peer nccl; do
    timeout 900 $T bench.py --gpus $N --steps 30 --warmup 5 --workload $W --exchange $X --no-cpu-baseline > gpurun_out/bench_${W}_${X}_n$N.json 2> gpurun_out/bench_${W}_${X}_n$N.err
    echo "$W $X N=$N exit $?"
    python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/bench_${W}_${X}_n$N.json") if l.startswith("{")][-1])
    print("  ms/step %.4f value %.0f e2e %.0f resident %.0f launches %d"%(d["ms_per_step"],d["value"],d["e2e"]["value"],d["e2e_resident"]["value"],d["gpu_launches"]), d["roofline"]["kernel_ms_per_step"])
except Exception as e:
    print("  failed", e); print(open("gpurun_out/bench_${W}_${X}_n$N.err").read()[-1500:])
PY
  done
done
