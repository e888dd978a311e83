#!/bin/bash
mkdir -p gpurun_out
python - <<'P' > gpurun_out/micro4.log 2>&1
import ctypes as C, sys
sys.path.insert(0, '.')
from lunar_lander_b200 import _abi
for code in (2001, 2021, 2041, 2081, 2002, 2022, 2042, 2082):
    tf, mhz = C.c_double(), C.c_double()
    rc = _abi.lib.ll_microbench_fma(0, code, 0.3, C.byref(tf), C.byref(mhz))
    hmma_per_s_per_smsp = tf.value * 1e12 / 2 / 1024 / (148 * 4)
    print(code, rc, "K FFMA per HMMA", (code // 10) % 10, "warps per sub-partition", code % 10, "TF/s", tf.value, "cycles per HMMA per sub-partition", 1.965e9 / hmma_per_s_per_smsp, flush=True)
P
cat gpurun_out/micro4.log
