#!/bin/bash
# exact kernel templated on the run-time activation pitch: full GPU suite, smoke, configs[0] bench line, exact mode at scale
mkdir -p gpurun_out
O=gpurun_out
( time timeout 1800 python -m pytest tests -q -m gpu ) > $O/j49_pytest.log 2>&1
echo "pytest rc=$?" >> $O/j49_pytest.log
tail -6 $O/j49_pytest.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3
timeout 600 python bench.py --workload c1 > $O/j49_bench_c1.json 2> $O/j49_bench_c1.err
cut -c1-220 $O/j49_bench_c1.json
run() { name=$1; shift
  timeout 600 python bench.py "$@" --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 0 > $O/j49_$name.log 2>&1
  python - <<P
import json
for l in open("$O/j49_$name.log"):
    if l.startswith("{"):
        d = json.loads(l); print("$name", "%.4g" % d["value"], "%.4f" % d["roofline"]["frac"], "%.2f ms" % d["ms_per_step"], d["config"]["kernel"])
P
}
run c1_f32 --workload c1 --precision 32 --exact
run c1_pop32768 --workload c1 --pop-per-gpu 32768
run c3_f64 --workload c3 --precision 64 --exact --pop-per-gpu 16384
run c4_f64 --workload c4 --precision 64 --exact --pop-per-gpu 2048
