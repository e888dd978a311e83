#!/bin/bash
mkdir -p gpurun_out
python scripts/mma_probe.py 32_128_32 256 50 > gpurun_out/probe_c4.log 2>&1; cat gpurun_out/probe_c4.log
python scripts/mma_probe.py 8_16_32_16_8 128 50 > gpurun_out/probe_c5.log 2>&1; tail -6 gpurun_out/probe_c5.log
python scripts/mma_probe.py 16_16 128 50 > gpurun_out/probe_c1.log 2>&1; tail -6 gpurun_out/probe_c1.log
timeout 900 python -m pytest tests -m gpu -x -q -k "fast" > gpurun_out/pytest_fast.log 2>&1; tail -5 gpurun_out/pytest_fast.log
run() { name=$1; shift
  python bench.py "$@" --steps 2 --warmup 1 --no-cpu-baseline --e2e-steps 0 > gpurun_out/f16_$name.log 2>&1
  python - <<P
import json
for l in open("gpurun_out/f16_$name.log"):
    if l.startswith("{"):
        d = json.loads(l); print("$name", d["value"], d["roofline"]["frac"], d["ms_per_step"], d["config"]["kernel"])
P
}
run c4_w4 --workload c4 --pop-per-gpu 32768
LL_MMA_WARPS_PER_SM=6 run c4_w6 --workload c4 --pop-per-gpu 32768
run c3L50 --workload c3 --layouts 50 --pop-per-gpu 32768
