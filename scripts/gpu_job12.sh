#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --workload c4 --pop-per-gpu 8192 --steps 1 --warmup 1 --no-cpu-baseline --e2e-steps 0"
$CMD > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:ll_episode_mma -s 1 -c 1 -o gpurun_out/prof_mma_f16 $CMD > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/ncu.log; tail -c 600 gpurun_out/plain.log
