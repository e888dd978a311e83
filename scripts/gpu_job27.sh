#!/bin/bash
# round-1 record run: bench lines for every BASELINE config, reference arm, ncu of the transposed kernel on configs[2],
# configs[4] end to end (10 000 generations with main_train_f, then headless nn_play of the written file)
mkdir -p gpurun_out
O=gpurun_out
timeout 300 python -m pytest tests/test_gpu_host.py -q -m gpu 2>&1 | tail -3
timeout 900 python bench.py > $O/j27_bench_c4.json 2> $O/j27_bench_c4.err
for w in c3 c2 c1; do timeout 600 python bench.py --workload $w > $O/j27_bench_$w.json 2> $O/j27_bench_$w.err; done
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/j27_bench_reference_c4.json 2> $O/j27_bench_reference_c4.err
for f in c4 c3 c2 c1 reference_c4; do cut -c1-330 $O/j27_bench_$f.json; done
CMD="python bench.py --workload c3 --pop-per-gpu 16384 --steps 2 --warmup 1 --no-cpu-baseline --e2e-steps 0"
$CMD > $O/j27_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/j27_launches_c3_pop16384.csv $CMD > $O/j27_ncu_list.log 2>&1
$CMD > $O/j27_plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:ll_episode_mmat -s 1 -c 1 -o $O/j27_prof_mmat_c3 $CMD > $O/j27_ncu_full.log 2>&1
tail -2 $O/j27_ncu_full.log
ncu -i $O/j27_prof_mmat_c3.ncu-rep --page raw --csv > $O/j27_prof_mmat_c3_raw.csv 2>/dev/null
# configs[4]
W=/tmp/c5run ; mkdir -p $W ; cp config.txt $W/
( cd $W && ( time /root/repo/lunar_lander_b200/host/bin/main_train_f --set NNConfig.hlayers=8,16,32,16,8 --set NNConfig.epochs=10000 --set NNConfig.save_interval=2500 > out.log 2> err.log ) 2> time.log
  tail -8 out.log; tail -1 err.log; tail -3 time.log; ls data )
cp $W/out.log $O/j27_c5_train.log; cp $W/err.log $O/j27_c5_train.err; cp $W/data/lunar_lander_last.txt $O/j27_c5_lunar_lander_last.txt
( cd $W && sed -i 's/^NNConfig.hlayers.*/NNConfig.hlayers = 8,16,32,16,8/' config.txt && PYTHONPATH=/root/repo timeout 600 python -m lunar_lander_b200.nn_play --seed 100 --episodes 30 ) > $O/j27_c5_nn_play.log 2>&1
tail -4 $O/j27_c5_nn_play.log
