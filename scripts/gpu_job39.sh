#!/bin/bash
# two accumulators (shared scaled small products), m16n8k8 for the first layer: parity of the fast kernels + c3, c2
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "fast or several_members or tensor" > gpurun_out/j39_tests.log 2>&1
tail -3 gpurun_out/j39_tests.log
for w in 15 14; do
  LL_MMA_WARPS_PER_SM=$w python bench.py --workload c3 --steps 10 --warmup 3 --no-cpu-baseline --e2e-steps 0 > gpurun_out/j39_c3_w$w.json 2> gpurun_out/j39_c3_w$w.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/j39_c3_w$w.json").read().strip().splitlines()[-1])
print("c3 warps $w", d["value"], d["roofline"]["frac"], d["ms_per_step"])
PY
done
python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline --e2e-steps 0 > gpurun_out/j39_c2.json 2>gpurun_out/j39_c2.err
python -c "
import json
d=json.loads(open('gpurun_out/j39_c2.json').read().strip().splitlines()[-1]); print('c2', d['value'], d['ms_per_step'])"
python bench.py --workload c3 --pop-per-gpu 131072 --layouts 50 --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 0 > gpurun_out/j39_c3x50.json 2>gpurun_out/j39_c3x50.err
python -c "
import json
d=json.loads(open('gpurun_out/j39_c3x50.json').read().strip().splitlines()[-1]); print('c3 topology x 50 layouts (mma kernel)', d['value'], d['roofline']['frac'], d['ms_per_step'], d['config']['kernel'])"
