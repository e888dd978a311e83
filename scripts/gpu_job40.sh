#!/bin/bash
# one fp32 accumulator tile for the three split products (unscaled lo parts): parity of the fast kernels + c4, c3, c2
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "fast or several_members or tensor" > gpurun_out/j40_tests.log 2>&1
tail -3 gpurun_out/j40_tests.log
show() { python - <<PY
import json
d=json.loads(open("gpurun_out/j40_$1.json").read().strip().splitlines()[-1])
print("$1", "%.4g" % d["value"], "%.4f" % d["roofline"]["frac"], "%.2f ms" % d["ms_per_step"], d["config"]["kernel"])
PY
}
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 0 > gpurun_out/j40_c4.json 2> gpurun_out/j40_c4.err; show c4
python bench.py --workload c3 --steps 10 --warmup 3 --no-cpu-baseline --e2e-steps 0 > gpurun_out/j40_c3.json 2> gpurun_out/j40_c3.err; show c3
python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline --e2e-steps 0 > gpurun_out/j40_c2.json 2> gpurun_out/j40_c2.err; show c2
python bench.py --workload c3 --pop-per-gpu 131072 --layouts 50 --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 0 > gpurun_out/j40_c3x50.json 2>gpurun_out/j40_c3x50.err; show c3x50
