#!/bin/bash
mkdir -p gpurun_out
python bench.py > gpurun_out/bench_c4_default.json 2> gpurun_out/bench_c4_default.err; tail -c 400 gpurun_out/bench_c4_default.err; cut -c1-600 gpurun_out/bench_c4_default.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_c4_reference.json 2>&1; cut -c1-400 gpurun_out/bench_c4_reference.json
CMD="python bench.py --pop-per-gpu 8192 --steps 2 --warmup 1 --no-cpu-baseline --e2e-steps 0"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c4_pop8192.csv $CMD > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/ncu2.log
