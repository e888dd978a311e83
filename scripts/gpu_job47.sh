#!/bin/bash
# ncu of the exact fp64 kernel at scale (16_16 x 50 layouts, 32768 members; 16_64_32 x 8, 16384 members)
mkdir -p gpurun_out
O=gpurun_out
CMD="python bench.py --workload c1 --pop-per-gpu 32768 --steps 1 --warmup 1 --no-cpu-baseline --e2e-steps 0"
$CMD > $O/j47_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:ll_episode_kernel -s 1 -c 1 -o $O/j47_prof_exact_c1 $CMD > $O/j47_ncu1.log 2>&1
ncu -i $O/j47_prof_exact_c1.ncu-rep --page raw --csv > $O/j47_prof_exact_c1_raw.csv 2>/dev/null
ncu -i $O/j47_prof_exact_c1.ncu-rep --page source --csv > $O/j47_prof_exact_c1_source.csv 2>/dev/null
rm -f $O/j47_prof_exact_c1.ncu-rep
tail -2 $O/j47_ncu1.log
CMD="python bench.py --workload c3 --precision 64 --exact --pop-per-gpu 16384 --steps 1 --warmup 1 --no-cpu-baseline --e2e-steps 0"
$CMD > $O/j47_plain2.log 2>&1 && ncu --set full --clock-control none -k regex:ll_episode_kernel -s 1 -c 1 -o $O/j47_prof_exact_c3 $CMD > $O/j47_ncu2.log 2>&1
ncu -i $O/j47_prof_exact_c3.ncu-rep --page raw --csv > $O/j47_prof_exact_c3_raw.csv 2>/dev/null
rm -f $O/j47_prof_exact_c3.ncu-rep
tail -2 $O/j47_ncu2.log
ls -la $O/j47*
