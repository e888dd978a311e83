#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
( time timeout 1500 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "fast or chunks or lane_split or generation_bit_exact" -x ) > $O/j28_pytest.log 2>&1
echo "pytest rc=$?" >> $O/j28_pytest.log
tail -6 $O/j28_pytest.log
run() { name=$1; shift
  timeout 600 python bench.py "$@" --steps 3 --warmup 2 --no-cpu-baseline --e2e-steps 0 > $O/j28_$name.log 2>&1
  python - <<P
import json
for l in open("$O/j28_$name.log"):
    if l.startswith("{"):
        d = json.loads(l); print("$name", "%.4g" % d["value"], "%.4f" % d["roofline"]["frac"], "%.2f ms" % d["ms_per_step"], d["config"]["kernel"])
P
}
