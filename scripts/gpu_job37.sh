#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
run() { name=$1; shift
  timeout 600 python bench.py "$@" --steps 3 --warmup 2 --no-cpu-baseline --e2e-steps 0 > $O/j37_$name.log 2>&1
  python - <<P
import json
for l in open("$O/j37_$name.log"):
    if l.startswith("{"):
        d = json.loads(l); print("$name", "%.4g" % d["value"], "%.4f" % d["roofline"]["frac"], "%.2f ms" % d["ms_per_step"], d["config"]["kernel"])
P
}
run c3_w14 --workload c3
LL_MMA_WARPS_PER_SM=13 run c3_w13 --workload c3
LL_MMA_WARPS_PER_SM=12 run c3_w12 --workload c3
