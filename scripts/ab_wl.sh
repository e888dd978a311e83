#!/bin/bash
# A/B of environment switches on one of the small workloads: scripts/ab_wl.sh c2 "LL_WARP_HALF=0" "LL_WARP_HALF=1" ...
WL=$1; shift
mkdir -p gpurun_out
n=0
for setting in "$@"; do
  n=$((n + 1))
  env $setting timeout 300 python bench.py --workload $WL --steps 20 --warmup 3 --auto-steps --no-cpu-baseline --no-extras --no-invariance --e2e-steps 3 \
      > gpurun_out/abwl_$n.json 2> gpurun_out/abwl_$n.err || { echo "[$setting] failed"; tail -5 gpurun_out/abwl_$n.err; }
  python - "$setting" gpurun_out/abwl_$n.json <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print("[%s]" % sys.argv[1], "value %.4e" % d["value"], "ms/step %.3f" % d["ms_per_step"], "K1 ms %.3f" % d["roofline"]["kernel_ms_per_launch"],
          "e2e %.4e" % d["e2e"]["value"], "steps/gen %.4e" % d["lander_steps_per_generation"], flush=True)
except Exception as e:
    print("[%s]" % sys.argv[1], "no result:", e, flush=True)
PY
done
