#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift
  timeout 600 python bench.py "$@" --steps 2 --warmup 1 --no-cpu-baseline --e2e-steps 0 > gpurun_out/j24_$name.log 2>&1
  python - <<P
import json
for l in open("gpurun_out/j24_$name.log"):
    if l.startswith("{"):
        d = json.loads(l); print("$name", "%.4g" % d["value"], "%.3f" % d["roofline"]["frac"], "%.2f ms" % d["ms_per_step"], d["config"]["kernel"])
P
}
LL_FAST_KERNEL=mmat run c4_mmat_nt2
LL_FAST_KERNEL=mmat LL_MMAT_TILES=1 run c4_mmat_nt1
LL_FAST_KERNEL=mma run c4_mma
LL_FAST_KERNEL=mmat run c3_mmat --workload c3
( time timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "fast" ) 2>&1 | tail -5
