#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift
  python bench.py "$@" --steps 1 --warmup 1 --no-cpu-baseline --e2e-steps 0 > gpurun_out/ex_$name.log 2>&1
  python - <<P
import json
for l in open("gpurun_out/ex_$name.log"):
    if l.startswith("{"):
        d = json.loads(l); print("$name", d["value"], d["roofline"]["frac"], d["ms_per_step"])
P
}
for w in 4 6; do for e in 9 1 2 3 4 5; do LL_MMA_WARPS_PER_SM=$w LL_MMA_EXP=$e run w${w}_exp$e --workload c4 --pop-per-gpu 4096; done; done
