#!/bin/bash
# microbench + current-state profile of the fast kernel
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
python - <<'P' > gpurun_out/micro.log 2>&1
import ctypes as C, sys
sys.path.insert(0, '.')
from lunar_lander_b200 import _abi
for code in (32, 322, 19, 191, 16):
    tf, mhz = C.c_double(), C.c_double()
    rc = _abi.lib.ll_microbench_fma(0, code, 0.5, C.byref(tf), C.byref(mhz))
    print(code, rc, tf.value, mhz.value, flush=True)
P
cat gpurun_out/micro.log
LL_FAST_PROF=1 python bench.py --workload c4 --pop-per-gpu 16384 --steps 2 --warmup 1 --no-cpu-baseline --e2e-steps 0 > gpurun_out/c4_prof.log 2>&1
tail -5 gpurun_out/c4_prof.log
LL_FAST_PROF=1 python bench.py --workload c3 --steps 2 --warmup 1 --no-cpu-baseline --e2e-steps 0 > gpurun_out/c3_prof.log 2>&1
tail -5 gpurun_out/c3_prof.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
tail -5 gpurun_out/pytest_gpu.log
python bench.py --workload c4 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/c4_full.log 2>&1
tail -3 gpurun_out/c4_full.log
