#!/bin/bash
# record run on the single-accumulator tensor-core kernels: full GPU suite, smoke, bench lines of every config, reference arm,
# ncu launch lists + one full capture each for configs[3] (per-GPU shard, pop 8192 sample) and configs[2] (pop 16384 sample)
mkdir -p gpurun_out
O=gpurun_out
( time timeout 1800 python -m pytest tests -q -m gpu ) > $O/j41_pytest.log 2>&1
echo "pytest rc=$?" >> $O/j41_pytest.log
tail -6 $O/j41_pytest.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3
timeout 900 python bench.py > $O/j41_bench_c4.json 2> $O/j41_bench_c4.err
for w in c3 c2 c1; do timeout 600 python bench.py --workload $w > $O/j41_bench_$w.json 2> $O/j41_bench_$w.err; done
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/j41_bench_reference_c4.json 2> $O/j41_bench_reference_c4.err
for f in c4 c3 c2 c1 reference_c4; do cut -c1-200 $O/j41_bench_$f.json; done
CMD="python bench.py --pop-per-gpu 8192 --steps 2 --warmup 1 --no-cpu-baseline --e2e-steps 0"
$CMD > $O/j41_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/j41_launches_c4_pop8192.csv $CMD > $O/j41_ncu_list.log 2>&1
$CMD > $O/j41_plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:ll_episode_mma -s 1 -c 1 -o $O/j41_prof_mma_c4 $CMD > $O/j41_ncu_full.log 2>&1
ncu -i $O/j41_prof_mma_c4.ncu-rep --page raw --csv > $O/j41_prof_mma_c4_raw.csv 2>/dev/null
rm -f $O/j41_prof_mma_c4.ncu-rep
tail -2 $O/j41_ncu_full.log
CMD="python bench.py --workload c3 --pop-per-gpu 16384 --steps 2 --warmup 1 --no-cpu-baseline --e2e-steps 0"
$CMD > $O/j41_plain3.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/j41_launches_c3_pop16384.csv $CMD > $O/j41_ncu_list3.log 2>&1
$CMD > $O/j41_plain4.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:ll_episode_mmat -s 1 -c 1 -o $O/j41_prof_mmat_c3 $CMD > $O/j41_ncu_full3.log 2>&1
ncu -i $O/j41_prof_mmat_c3.ncu-rep --page raw --csv > $O/j41_prof_mmat_c3_raw.csv 2>/dev/null
rm -f $O/j41_prof_mmat_c3.ncu-rep
tail -2 $O/j41_ncu_full3.log
