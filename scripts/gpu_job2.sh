#!/bin/bash
set -x
mkdir -p gpurun_out
python scripts/mma_probe.py 32_128_32 256 50 > gpurun_out/probe_c4.log 2>&1; cat gpurun_out/probe_c4.log
python scripts/mma_probe.py 8_16_32_16_8 128 50 > gpurun_out/probe_c5.log 2>&1; tail -9 gpurun_out/probe_c5.log
timeout 900 python -m pytest tests -m gpu -x -q -k "fast" > gpurun_out/pytest_fast.log 2>&1; tail -5 gpurun_out/pytest_fast.log
for cfg in "1 3 0" "2 3 0" "1 2 0" "2 2 0" "1 3 4" "1 3 5"; do
  set -- $cfg
  export LL_MMA_MT=$1 LL_MMA_TERMS=$2
  if [ "$3" != "0" ]; then export LL_MMA_WARPS_PER_SM=$3; else unset LL_MMA_WARPS_PER_SM; fi
  python bench.py --workload c4 --pop-per-gpu 32768 --steps 2 --warmup 1 --no-cpu-baseline --e2e-steps 0 > gpurun_out/c4_mma_$1_$2_$3.log 2>&1
  python - <<P
import json
for l in open("gpurun_out/c4_mma_$1_$2_$3.log"):
    if l.startswith("{"):
        d = json.loads(l); print("MT=$1 TERMS=$2 WPSM=$3", d["value"], d["roofline"]["frac"], d["config"]["kernel"], d["clocks"])
P
done
