#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
run() { name=$1; shift
  timeout 600 python bench.py "$@" --steps 2 --warmup 1 --no-cpu-baseline --e2e-steps 0 > $O/j29_$name.log 2>&1
  python - <<P
import json
for l in open("$O/j29_$name.log"):
    if l.startswith("{"):
        d = json.loads(l); print("$name", "%.4g" % d["value"], "%.4f" % d["roofline"]["frac"], "%.2f ms" % d["ms_per_step"], d["config"]["kernel"])
P
}
run c4_mma_w6
LL_MMA_WARPS_PER_SM=4 run c4_mma_w4
LL_FAST_KERNEL=mmat LL_MMAT_TILES=1 run c4_mmat1_w6
LL_FAST_KERNEL=mmat LL_MMAT_TILES=1 LL_MMA_WARPS_PER_SM=5 run c4_mmat1_w5
