#!/bin/bash
# C++ host mirror vs the compiled reference main_train; timing of BASELINE configs[4] (8_16_32_16_8) generations
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests/test_gpu_host.py -x -q -m gpu ) > gpurun_out/j20_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/j20_pytest.log
tail -25 gpurun_out/j20_pytest.log
W=/tmp/c5 ; mkdir -p $W ; cp config.txt $W/
for b in d f; do
  ( cd $W && rm -rf data && time timeout 600 /root/repo/lunar_lander_b200/host/bin/main_train_$b --set NNConfig.hlayers=8,16,32,16,8 --set NNConfig.epochs=300 > /root/repo/gpurun_out/j20_c5_$b.log 2> /root/repo/gpurun_out/j20_c5_$b.err )
  tail -8 gpurun_out/j20_c5_$b.log; tail -4 gpurun_out/j20_c5_$b.err
done
( cd $W && rm -rf data && time timeout 600 /root/repo/oracle/_ref/main_train_d > /dev/null 2>&1 ) 2>&1 | tail -4
