#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift
  python bench.py "$@" --steps 2 --warmup 1 --no-cpu-baseline --e2e-steps 0 > gpurun_out/t_$name.log 2>&1
  python - <<P
import json
for l in open("gpurun_out/t_$name.log"):
    if l.startswith("{"):
        d = json.loads(l); print("$name", d["value"], d["roofline"]["frac"], d["ms_per_step"], d["config"]["kernel"])
P
}
run c3_ffma --workload c3
LL_FAST_KERNEL=mma run c3_mma --workload c3
LL_FAST_KERNEL=mma run c3_L4_mma --workload c3 --layouts 4
run c3_L4_ffma --workload c3 --layouts 4
LL_FAST_KERNEL=mma run c2_mma --workload c2
run c2_default --workload c2
