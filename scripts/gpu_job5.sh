#!/bin/bash
mkdir -p gpurun_out
run() { # name, workload args..., env via VARS
  name=$1; shift
  python bench.py "$@" --steps 2 --warmup 1 --no-cpu-baseline --e2e-steps 0 > gpurun_out/sw_$name.log 2>&1
  python - <<P
import json
for l in open("gpurun_out/sw_$name.log"):
    if l.startswith("{"):
        d = json.loads(l); print("$name", d["value"], d["roofline"]["frac"], d["ms_per_step"])
P
}
for w in 1 2 3 4 6; do LL_MMA_WARPS_PER_SM=$w run c4_w$w --workload c4 --pop-per-gpu 16384; done
for w in 2 4 6 8 10 12; do LL_MMA_WARPS_PER_SM=$w run c3L50_w$w --workload c3 --layouts 50 --pop-per-gpu 32768; done
