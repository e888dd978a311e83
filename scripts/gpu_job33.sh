#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
( time timeout 900 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "several_members" ) > $O/j33_pytest.log 2>&1
echo "pytest rc=$?" >> $O/j33_pytest.log
tail -12 $O/j33_pytest.log
run() { name=$1; shift
  timeout 600 python bench.py "$@" --steps 3 --warmup 2 --no-cpu-baseline --e2e-steps 0 > $O/j33_$name.log 2>&1
  python - <<P
import json
for l in open("$O/j33_$name.log"):
    if l.startswith("{"):
        d = json.loads(l); print("$name", "%.4g" % d["value"], "%.4f" % d["roofline"]["frac"], "%.2f ms" % d["ms_per_step"], d["config"]["kernel"])
P
}
for nm in 0 2 3 4; do LL_MMAT_MEMBERS=$nm run c3_m$nm --workload c3; done
for nm in 0 4; do LL_MMAT_MEMBERS=$nm run c2_m$nm --workload c2; done
