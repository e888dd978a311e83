#!/bin/bash
mkdir -p gpurun_out
python scripts/mma_probe.py 32_128_32 256 50 > gpurun_out/probe_c4.log 2>&1; tail -5 gpurun_out/probe_c4.log
python scripts/mma_probe.py 8_16_8 128 33 > gpurun_out/probe_c2.log 2>&1; tail -5 gpurun_out/probe_c2.log
run() { name=$1; shift
  python bench.py "$@" --no-cpu-baseline --e2e-steps 0 > gpurun_out/f16b_$name.log 2>&1
  python - <<P
import json
for l in open("gpurun_out/f16b_$name.log"):
    if l.startswith("{"):
        d = json.loads(l); print("$name", d["value"], d["roofline"]["frac"], d["ms_per_step"], d["roofline"]["kernel_ms_per_launch"], d["lander_steps_per_generation"])
P
}
LL_MMA_EXP=9 run exp9 --workload c4 --pop-per-gpu 4736 --steps 1 --warmup 1
run c4 --workload c4 --pop-per-gpu 32768 --steps 2 --warmup 1
