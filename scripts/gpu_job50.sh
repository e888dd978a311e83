#!/bin/bash
# configs[4] end to end on the final kernels: 10 000 generations of 8_16_32_16_8 with main_train_f, then headless nn_play of the file
mkdir -p gpurun_out
O=gpurun_out
W=/tmp/c5run ; mkdir -p $W ; cp config.txt $W/
( cd $W && ( time /root/repo/lunar_lander_b200/host/bin/main_train_f --set NNConfig.hlayers=8,16,32,16,8 --set NNConfig.epochs=10000 --set NNConfig.save_interval=2500 > out.log 2> err.log ) 2> time.log
  tail -8 out.log; tail -1 err.log; tail -3 time.log; ls data )
cp $W/out.log $O/j50_c5_train.log; cp $W/err.log $O/j50_c5_train.err; cp $W/time.log $O/j50_c5_time.log
( cd $W && sed -i 's/^NNConfig.hlayers.*/NNConfig.hlayers = 8,16,32,16,8/' config.txt && PYTHONPATH=/root/repo timeout 600 python -m lunar_lander_b200.nn_play --seed 100 --episodes 30 ) > $O/j50_c5_nn_play.log 2>&1
tail -4 $O/j50_c5_nn_play.log
# the same in bit-exact fp64 (main_train_d), 2000 generations
W=/tmp/c5d ; mkdir -p $W ; cp config.txt $W/
( cd $W && ( time /root/repo/lunar_lander_b200/host/bin/main_train_d --set NNConfig.hlayers=8,16,32,16,8 --set NNConfig.epochs=2000 --set NNConfig.save_interval=1000 > out.log 2> err.log ) 2> time.log
  tail -4 out.log; tail -3 time.log )
cp $W/out.log $O/j50_c5d_train.log; cp $W/time.log $O/j50_c5d_time.log
