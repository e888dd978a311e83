#!/bin/bash
# latency check of the M-side kernel on a latency-bound workload (configs[1] forced onto it; before the single accumulator: 6.0e8)
mkdir -p gpurun_out
O=gpurun_out
run() { name=$1; shift
  timeout 600 python bench.py "$@" --steps 10 --warmup 3 --no-cpu-baseline --e2e-steps 0 > $O/j51_$name.log 2>&1
  python - <<P
import json
for l in open("$O/j51_$name.log"):
    if l.startswith("{"):
        d = json.loads(l); print("$name", "%.4g" % d["value"], "%.4f" % d["roofline"]["frac"], "%.3f ms" % d["ms_per_step"], d["config"]["kernel"])
P
}
LL_FAST_KERNEL=mma run c2_mma --workload c2
run c2_mmat --workload c2
LL_FAST_KERNEL=mma run c2_100x50_mma --workload c2 --pop-per-gpu 100 --layouts 50
run c2_100x50 --workload c2 --pop-per-gpu 100 --layouts 50
