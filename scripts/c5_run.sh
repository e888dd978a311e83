#!/bin/bash
# BASELINE.json configs[4] end to end: main_train_f trains the default population of the 8_16_32_16_8 network for 10 000 generations,
# writes data/lunar_lander_last.txt, and the headless nn_play loop flies 30 unseen layouts with the saved best member.
# Outputs: gpurun_out/c5_main_train_f_10000gen.log (head + tail), gpurun_out/c5_nn_play_30_episodes.log
set -u
mkdir -p gpurun_out
d=$(mktemp -d); cp config.txt $d/; mkdir -p $d/data
( cd $d && time timeout 900 /root/repo/lunar_lander_b200/host/bin/main_train_f --config config.txt --set NNConfig.hlayers=8,16,32,16,8 --set NNConfig.epochs=10000 \
      --set NNConfig.save_interval=2500 ) > $d/train.log 2>&1
echo "main_train_f rc=$?"
( head -20 $d/train.log; echo "[...]"; tail -14 $d/train.log ) > gpurun_out/c5_main_train_f_10000gen.log
tail -6 gpurun_out/c5_main_train_f_10000gen.log
( cd $d && PYTHONPATH=/root/repo timeout 600 python -m lunar_lander_b200.nn_play --config config.txt --seed 100 --episodes 30 ) > gpurun_out/c5_nn_play_30_episodes.log 2>&1
tail -1 gpurun_out/c5_nn_play_30_episodes.log
