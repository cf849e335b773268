#!/bin/bash
# the driver's scaling launch at N GPUs: own arm (default workload) and the reference arm, both under torchrun
N=${N:-4}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name --format=csv,noheader | sort | uniq -c > gpurun_out/scale_gpu_n$N.txt; nproc >> gpurun_out/scale_gpu_n$N.txt; free -g | head -2 >> gpurun_out/scale_gpu_n$N.txt
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 1200 $T bench.py --gpus $N --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/scale_own_n$N.json 2> gpurun_out/scale_own_n$N.err
echo "own N=$N exit $?"; cat gpurun_out/scale_own_n$N.json; tail -5 gpurun_out/scale_own_n$N.err
timeout 900 $T bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/scale_ref_n$N.json 2> gpurun_out/scale_ref_n$N.err
echo "ref N=$N exit $?"; cat gpurun_out/scale_ref_n$N.json; tail -5 gpurun_out/scale_ref_n$N.err
cat gpurun_out/scale_gpu_n$N.txt
