#!/bin/bash
# configs[0] with the neurons of a slot dealt to Q lanes, now that the tanh chains run side by side
mkdir -p gpurun_out
O=gpurun_out
run() { name=$1; shift
  timeout 600 python bench.py "$@" --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 0 > $O/j44_$name.log 2>&1
  python - <<P
import json
for l in open("$O/j44_$name.log"):
    if l.startswith("{"):
        d = json.loads(l); print("$name", "%.4g" % d["value"], "%.4f" % d["roofline"]["frac"], "%.2f ms" % d["ms_per_step"], d["config"]["kernel"])
P
}
for q in 1 2 4 8; do LL_EXACT_Q=$q run c1_q$q --workload c1; done
