#!/bin/bash
# N-GPU bench (own arm) through torchrun: c4 (one camera per rank batch) and c5 (one camera per instance), gather to rank 0 inside the step
N=${N:-2}
mkdir -p gpurun_out
for W in c4 c5; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 --workload $W --no-cpu-baseline > gpurun_out/bench_${W}_n$N.json 2> gpurun_out/bench_${W}_n$N.err
  echo "$W N=$N exit $?"; cat gpurun_out/bench_${W}_n$N.json; tail -5 gpurun_out/bench_${W}_n$N.err
done
timeout 600 python bench.py --steps 20 --warmup 5 --workload c5 --no-cpu-baseline > gpurun_out/bench_c5_n1.json 2> gpurun_out/bench_c5_n1.err; echo "c5 N=1 exit $?"; cat gpurun_out/bench_c5_n1.json; tail -3 gpurun_out/bench_c5_n1.err
