#!/bin/bash
# re-entry check of HEAD: GPU tests, smoke, default bench lines
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/j19_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/j19_pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/j19_smoke.log 2>&1
echo "smoke rc=$?" >> gpurun_out/j19_smoke.log
timeout 600 python bench.py > gpurun_out/j19_bench_c4.json 2> gpurun_out/j19_bench_c4.err
for w in c3 c2; do
  timeout 300 python bench.py --workload $w > gpurun_out/j19_bench_$w.json 2> gpurun_out/j19_bench_$w.err
done
tail -3 gpurun_out/j19_pytest.log; tail -3 gpurun_out/j19_smoke.log
