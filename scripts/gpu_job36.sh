#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
( time timeout 1500 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "not fast_mode and not tracks and not config" ) > $O/j36_pytest.log 2>&1
echo "pytest rc=$?" >> $O/j36_pytest.log
tail -8 $O/j36_pytest.log
run() { name=$1; shift
  timeout 600 python bench.py "$@" --steps 2 --warmup 1 --no-cpu-baseline --e2e-steps 0 > $O/j36_$name.log 2>&1
  python - <<P
import json
for l in open("$O/j36_$name.log"):
    if l.startswith("{"):
        d = json.loads(l); print("$name", "%.4g" % d["value"], "%.4f" % d["roofline"]["frac"], "%.2f ms" % d["ms_per_step"], d["config"]["kernel"])
P
}

LL_EXACT_Q=1 run c3_f64_q1 --workload c3 --precision 64 --exact --pop-per-gpu 16384
run c3_f64 --workload c3 --precision 64 --exact --pop-per-gpu 16384
