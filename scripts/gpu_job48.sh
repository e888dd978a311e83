#!/bin/bash
# exact kernel with Q > 1: compact activation columns, bank-tiling row pitch: bit-exact parity + configs[0], 16_64_32 fp64
mkdir -p gpurun_out
O=gpurun_out
( time timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_host.py tests/test_gpu_play.py -x -q -m gpu -k "not fast" ) > $O/j48_pytest.log 2>&1
echo "pytest rc=$?" >> $O/j48_pytest.log
tail -4 $O/j48_pytest.log
run() { name=$1; shift
  timeout 600 python bench.py "$@" --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 0 > $O/j48_$name.log 2>&1
  python - <<P
import json
for l in open("$O/j48_$name.log"):
    if l.startswith("{"):
        d = json.loads(l); print("$name", "%.4g" % d["value"], "%.4f" % d["roofline"]["frac"], "%.2f ms" % d["ms_per_step"], d["config"]["kernel"])
P
}
run c1 --workload c1
LL_EXACT_Q=8 run c1_q8 --workload c1
LL_EXACT_Q=2 run c1_q2 --workload c1
run c1_f32 --workload c1 --precision 32 --exact
run c3_f64 --workload c3 --precision 64 --exact --pop-per-gpu 16384
run c3_f32_exact --workload c3 --precision 32 --exact --pop-per-gpu 16384
