#!/bin/bash
# Multi-GPU evidence in one call (run under `gpurun --gpus N -- scripts/multi_gpu_job.sh N`):
#   the 2-GPU tests of the C++ trainer, the weak- and strong-scaling bench lines at N ranks (torchrun, one rank per GPU), and at
#   N = 8 the C++ trainer on the 2^20-member configs[3] population (GpuCfg.gpus = 8).  Outputs in gpurun_out/.
set -u
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_host.py -m gpu -x -q -k two_gpus 2>&1 | tail -3
bench() { # name port args...
  local name=$1 port=$2; shift 2
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port bench.py --gpus $N "$@" \
      > gpurun_out/$name.json 2> gpurun_out/$name.err
  echo "$name rc=$?"
  python - gpurun_out/$name.json <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("  value %.4e e2e %.4e ms/step %.1f gen/s %.3f scaling %s K1 ms by rank %s invariance %s" % (
        d["value"], d["e2e"]["value"], d["ms_per_step"], d["generations_per_sec"], d["scaling"],
        [round(x, 1) for x in d["roofline"]["kernel_ms_per_launch_by_rank"]], ((d.get("invariance") or {}).get("sha256") or "")[:16]))
except Exception as e:
    print("  no result:", e)
PY
}
bench bench_c4_${N}gpu_weak 29511 --steps 3 --warmup 3 --no-extras
bench bench_c4_${N}gpu_strong 29512 --steps 3 --warmup 3 --no-extras --no-invariance --scaling strong
if [ "$N" -ge 8 ]; then
  d=$(mktemp -d); cp config.txt $d/; mkdir -p $d/data
  ( cd $d && time timeout 600 /root/repo/lunar_lander_b200/host/bin/main_train_f --config config.txt --set GpuCfg.gpus=8 --set NNConfig.hlayers=32,128,32 \
      --set NNConfig.population_size=1048576 --set NNConfig.epochs=2 --set NNConfig.save_interval=2 ) > gpurun_out/cpp_main_train_f_8gpu_1M.log 2>&1
  echo "main_train_f 8 GPUs rc=$?"; tail -8 gpurun_out/cpp_main_train_f_8gpu_1M.log
fi
