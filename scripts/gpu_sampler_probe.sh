#!/bin/bash
# per-launch time and DRAM bytes of the track sampler over consecutive frames (cursor walks through the keys)
mkdir -p gpurun_out
B="python bench.py --steps 12 --warmup 3 --no-cpu-baseline"
timeout 600 $B > gpurun_out/probe_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,l1tex__data_pipe_lsu_wavefronts.sum --clock-control none -k regex:k_sample_tracks -c 40 --csv --log-file gpurun_out/sampler_probe.csv $B > gpurun_out/probe_ncu.log 2>&1
tail -3 gpurun_out/probe_ncu.log
