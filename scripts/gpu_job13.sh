#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -15 gpurun_out/pytest_gpu.log
python bench.py --workload c4 --pop-per-gpu 32768 --steps 2 --warmup 1 --no-cpu-baseline --e2e-steps 0 > gpurun_out/c4_lpt.log 2>&1
python - <<P
import json
for l in open("gpurun_out/c4_lpt.log"):
    if l.startswith("{"):
        d = json.loads(l); print("c4 lpt", d["value"], d["roofline"]["frac"], d["ms_per_step"], d["config"]["kernel"])
P
