#!/bin/bash
# A/B of environment switches on the benchmark workload (c4 shard), device-timed, no CPU legs:
#   scripts/ab_env.sh <pop_per_gpu> "LL_TC05_CUT=16" "LL_TC05_RESUME=mma LL_TC05_CUT=8" ...   -> gpurun_out/ab_<n>.json + one summary line each
POP=${1:-32768}; shift
mkdir -p gpurun_out
n=0
for setting in "$@"; do
  n=$((n + 1))
  env $setting timeout 600 python bench.py --pop-per-gpu $POP --steps 3 --warmup 3 --no-cpu-baseline --no-extras --no-invariance --e2e-steps 1 \
      > gpurun_out/ab_$n.json 2> gpurun_out/ab_$n.err || { echo "[$setting] failed"; tail -5 gpurun_out/ab_$n.err; }
  python - "$setting" gpurun_out/ab_$n.json <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print("[%s]" % sys.argv[1], "value %.4e" % d["value"], "ms/step %.2f" % d["ms_per_step"], "K1 ms %.2f" % d["roofline"]["kernel_ms_per_launch"],
          "frac %.3f" % d["roofline"]["frac"], "steps/gen %.4e" % d["lander_steps_per_generation"], flush=True)
except Exception as e:
    print("[%s]" % sys.argv[1], "no result:", e, flush=True)
PY
done
