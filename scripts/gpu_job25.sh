#!/bin/bash
# exact kernel with Q lanes per episode: bit-exact tests, then c1 timing per Q
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "not fast and not config" ) > gpurun_out/j25_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/j25_pytest.log
tail -8 gpurun_out/j25_pytest.log
run() { name=$1; shift
  timeout 600 python bench.py "$@" --steps 3 --warmup 2 --no-cpu-baseline --e2e-steps 0 > gpurun_out/j25_$name.log 2>&1
  python - <<P
import json
for l in open("gpurun_out/j25_$name.log"):
    if l.startswith("{"):
        d = json.loads(l); print("$name", "%.4g" % d["value"], "%.4f" % d["roofline"]["frac"], "%.2f ms" % d["ms_per_step"], d["config"]["kernel"])
P
}
for q in 1 2 4; do LL_EXACT_Q=$q run c1_q$q --workload c1; done
run c1_auto --workload c1
for q in 1 4 8; do LL_EXACT_Q=$q run c5d_q$q --workload c1 --pop-per-gpu 100; done
run c1_big --workload c1 --pop-per-gpu 32768
run c4_f64 --workload c4 --precision 64 --exact --pop-per-gpu 8192
run c4_f32x --workload c4 --exact --pop-per-gpu 8192
