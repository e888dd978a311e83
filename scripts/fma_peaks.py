#!/usr/bin/env python
"""FMA-pipe microbenchmarks of ll_microbench_fma (FFMA, packed FFMA2, DFMA, legacy f16 HMMA) on device 0."""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from lunar_lander_b200 import _abi  # noqa: E402

for code in (32, 322, 64, 197):
    tf, mhz = C.c_double(), C.c_double()
    rc = _abi.lib.ll_microbench_fma(0, code, 0.5, C.byref(tf), C.byref(mhz))
    print(f"code {code}: rc {rc}, {tf.value:.2f} TFLOP/s, sm clock estimate {mhz.value:.0f} MHz", flush=True)
