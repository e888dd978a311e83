#!/bin/bash
# transposed tensor-core kernel (slots on N): parity tests, then timing against the M-side kernel on c3 / c4 / c2
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "fast" ) > gpurun_out/j23_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/j23_pytest.log
tail -15 gpurun_out/j23_pytest.log
run() { name=$1; shift
  timeout 600 python bench.py "$@" --steps 2 --warmup 1 --no-cpu-baseline --e2e-steps 0 > gpurun_out/j23_$name.log 2>&1
  python - <<P
import json
for l in open("gpurun_out/j23_$name.log"):
    if l.startswith("{"):
        d = json.loads(l); print("$name", "%.4g" % d["value"], "%.3f" % d["roofline"]["frac"], "%.2f ms" % d["ms_per_step"], d["config"]["kernel"])
P
}
LL_FAST_KERNEL=mmat run c3_mmat --workload c3
LL_FAST_KERNEL=mma run c3_mma --workload c3
LL_FAST_KERNEL=mmat run c2_mmat --workload c2
LL_FAST_KERNEL=mma run c2_mma --workload c2
LL_FAST_KERNEL=mmat run c4_mmat
LL_FAST_KERNEL=mmat LL_MMAT_TILES=1 run c4_mmat_nt1
LL_FAST_KERNEL=mmat LL_MMAT_TILES=2 run c3_mmat_nt2 --workload c3
