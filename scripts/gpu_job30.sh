#!/bin/bash
# full GPU suite + smoke with the final kernels; default bench line; 
mkdir -p gpurun_out
O=gpurun_out
( time timeout 1800 python -m pytest tests -q -m gpu ) > $O/j30_pytest.log 2>&1
echo "pytest rc=$?" >> $O/j30_pytest.log
tail -6 $O/j30_pytest.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3
timeout 900 python bench.py > $O/j30_bench_c4.json 2> $O/j30_bench_c4.err
timeout 600 python bench.py --workload c3 > $O/j30_bench_c3.json 2> $O/j30_bench_c3.err
timeout 600 python bench.py --workload c1 > $O/j30_bench_c1.json 2> $O/j30_bench_c1.err
for f in c4 c3 c1; do cut -c1-200 $O/j30_bench_$f.json; done
CMD="python bench.py --pop-per-gpu 8192 --steps 2 --warmup 1 --no-cpu-baseline --e2e-steps 0"
$CMD > $O/j30_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/j30_launches_c4_pop8192.csv $CMD > $O/j30_ncu_list.log 2>&1
$CMD > $O/j30_plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:ll_episode_mma -s 1 -c 1 -o $O/j30_prof_mma_c4 $CMD > $O/j30_ncu_full.log 2>&1
ncu -i $O/j30_prof_mma_c4.ncu-rep --page raw --csv > $O/j30_prof_mma_c4_raw.csv 2>/dev/null
rm -f $O/j30_prof_mma_c4.ncu-rep
tail -2 $O/j30_ncu_full.log
