#!/bin/bash
# A/B of the fast-mode episode kernels on the benchmark workload (c4 shard): LL_FAST_KERNEL = mma | tc05, device-timed, no CPU legs.
# usage: scripts/bench_kernels.sh [pop_per_gpu] [kernels...]   -> gpurun_out/ab_<kernel>.json
POP=${1:-131072}; shift
KERNELS=${@:-"mma tc05"}
mkdir -p gpurun_out
for k in $KERNELS; do
  LL_FAST_KERNEL=$k timeout 600 python bench.py --pop-per-gpu $POP --steps 3 --warmup 3 --no-cpu-baseline --no-extras --no-invariance --e2e-steps 1 \
      > gpurun_out/ab_$k.json 2> gpurun_out/ab_$k.err || { echo "$k failed"; tail -5 gpurun_out/ab_$k.err; }
  python - "$k" <<'PY'
import json, sys
k = sys.argv[1]
try:
    d = json.loads(open(f"gpurun_out/ab_{k}.json").read().strip().splitlines()[-1])
    print(k, "value %.4e" % d["value"], "ms/step %.1f" % d["ms_per_step"], "frac %.3f" % d["roofline"]["frac"], d["roofline"]["kernel"], "steps/gen %.4e" % d["lander_steps_per_generation"])
except Exception as e:
    print(k, "no result:", e)
PY
done
