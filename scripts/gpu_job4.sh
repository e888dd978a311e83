#!/bin/bash
set -x
mkdir -p gpurun_out
python - <<'P' > gpurun_out/micro2.log 2>&1
import ctypes as C, sys
sys.path.insert(0, '.')
from lunar_lander_b200 import _abi
for code in (19, 191, 194, 195):
    tf, mhz = C.c_double(), C.c_double()
    rc = _abi.lib.ll_microbench_fma(0, code, 0.5, C.byref(tf), C.byref(mhz))
    warps = {19: 64, 191: 16, 194: 4, 195: 4}[code]
    hmma_per_s_per_smsp = tf.value * 1e12 / 2 / 1024 / (148 * 4)
    print(code, rc, "TF/s", tf.value, "cycles per HMMA per sub-partition @1965MHz", 1.965e9 / hmma_per_s_per_smsp, flush=True)
P
cat gpurun_out/micro2.log
run() { # name, env...
  name=$1; shift
  env "$@" python bench.py --workload c4 --pop-per-gpu 32768 --steps 2 --warmup 1 --no-cpu-baseline --e2e-steps 0 > gpurun_out/c4_$name.log 2>&1
  python - <<P
import json
for l in open("gpurun_out/c4_$name.log"):
    if l.startswith("{"):
        d = json.loads(l); print("$name", d["value"], d["roofline"]["frac"], d["config"]["kernel"])
P
}
run v2w6 LL_MMA_MT=1
run v2w4 LL_MMA_MT=1 LL_MMA_WARPS_PER_SM=4
python scripts/mma_probe.py 32_128_32 256 50 > gpurun_out/probe_c4.log 2>&1; tail -6 gpurun_out/probe_c4.log
