#!/bin/bash
set -x
mkdir -p gpurun_out
run() { # name, env...
  name=$1; shift
  env "$@" python bench.py --workload c4 --pop-per-gpu 32768 --steps 2 --warmup 1 --no-cpu-baseline --e2e-steps 0 > gpurun_out/c4_$name.log 2>&1
  python - <<P
import json
for l in open("gpurun_out/c4_$name.log"):
    if l.startswith("{"):
        d = json.loads(l); print("$name", d["value"], d["roofline"]["frac"], d["config"]["kernel"])
P
}
run w6 LL_MMA_MT=1
run w4 LL_MMA_MT=1 LL_MMA_WARPS_PER_SM=4
run w5 LL_MMA_MT=1 LL_MMA_WARPS_PER_SM=5
run mt2w6 LL_MMA_MT=2
run mt2w4 LL_MMA_MT=2 LL_MMA_WARPS_PER_SM=4
CMD="python bench.py --workload c4 --pop-per-gpu 2048 --steps 1 --warmup 1 --no-cpu-baseline --e2e-steps 0"
$CMD > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:ll_episode_mma -s 1 -c 1 -o gpurun_out/prof_mma_v1 $CMD > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/ncu.log
