#!/bin/bash
mkdir -p gpurun_out
python - <<'P' > gpurun_out/micro3.log 2>&1
import ctypes as C, sys
sys.path.insert(0, '.')
from lunar_lander_b200 import _abi
for code in (194, 196, 198, 197, 16):
    tf, mhz = C.c_double(), C.c_double()
    rc = _abi.lib.ll_microbench_fma(0, code, 0.5, C.byref(tf), C.byref(mhz))
    mac = 2048 if code in (197, 16) else 1024
    hmma_per_s_per_smsp = tf.value * 1e12 / 2 / mac / (148 * 4)
    print(code, rc, "TF/s", tf.value, "cycles per HMMA per sub-partition @1965MHz", 1.965e9 / hmma_per_s_per_smsp, flush=True)
P
cat gpurun_out/micro3.log
