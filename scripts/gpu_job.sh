#!/bin/bash
# One parametrised GPU job (replaces the one-off gpu_job<N>.sh of round 1).  Run under gpurun:
#   gpurun -- scripts/gpu_job.sh tests            # pytest -m gpu
#   gpurun -- scripts/gpu_job.sh bench [args]     # python bench.py [args]               -> gpurun_out/bench.json
#   gpurun -- scripts/gpu_job.sh ab [pop] [k...]  # fast-mode kernel A/B (bench_kernels.sh)
#   gpurun -- scripts/gpu_job.sh launches <cmd>   # ncu launch list of <cmd>             -> gpurun_out/launches.csv
#   gpurun -- scripts/gpu_job.sh ncu <regex> <cmd># ncu --set full of the kernels matching -> gpurun_out/prof.ncu-rep
set -u
mkdir -p gpurun_out
what=${1:-tests}; shift || true
case "$what" in
  tests)    timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tee gpurun_out/pytest_gpu.log | tail -15 ;;
  bench)    timeout 900 python bench.py "$@" > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "rc=$?"; tail -3 gpurun_out/bench.err ;;
  ab)       scripts/bench_kernels.sh "$@" ;;
  launches) "$@" > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv "$@" > gpurun_out/ncu.log 2>&1; tail -2 gpurun_out/plain.log ;;
  ncu)      re=$1; shift; "$@" > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"$re" -c ${NCU_COUNT:-2} -o gpurun_out/${NCU_OUT:-prof} "$@" > gpurun_out/ncu.log 2>&1; tail -2 gpurun_out/ncu.log ;;
  *) echo "unknown job $what"; exit 2 ;;
esac
