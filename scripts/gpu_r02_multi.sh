#!/bin/bash
# usage: gpurun --gpus N -- 'N=<n> bash scripts/gpu_r02_multi.sh'
mkdir -p gpurun_out
N=${N:-2}
nvidia-smi --query-gpu=index,name --format=csv,noheader > gpurun_out/gpus_n$N.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 900 $TR tests/mgpu_peer_check.py > gpurun_out/peer_check_n$N.log 2>&1; echo "peer_check exit $?"; grep -E "identical|fell back|Error|error|assert" gpurun_out/peer_check_n$N.log | head -12
run() { n=$1; shift; timeout 900 env $ENVV $TR bench.py --gpus $N --no-cpu-baseline "$@" > gpurun_out/${n}_n$N.json 2> gpurun_out/${n}_n$N.err; echo "$n exit $?"; python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/${n}_n$N.json") if l.startswith("{")][-1])
    print("  ms/step %.4f value %.0f e2e %.0f exch %s"%(d["ms_per_step"],d["value"],d["e2e"]["value"],d.get("exchange_check")), {k:round(v,4) for k,v in d["roofline"]["kernel_ms_per_step"].items()})
except Exception as e:
    print("  failed", e); print(open("gpurun_out/${n}_n$N.err").read()[-1500:])
PY
}
run bench_c4 --steps 100 --warmup 5
ENVV="HGB_ASYNC_POST=0" run bench_c4_sync --steps 100 --warmup 5
run bench_c5_strong --workload c5 --strong --steps 100 --warmup 5
run bench_c5_weak --workload c5 --steps 100 --warmup 5
if [ "$EXTRA" = "1" ]; then
  run bench_c4_nccl --exchange nccl --steps 50 --warmup 5
fi
ls gpurun_out | head -40
