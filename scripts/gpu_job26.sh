#!/bin/bash
# full GPU suite after the kernel changes; small-population timing of main_train_f with / without layout chunks
mkdir -p gpurun_out
( time timeout 1800 python -m pytest tests -q -m gpu ) > gpurun_out/j26_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/j26_pytest.log
tail -12 gpurun_out/j26_pytest.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -4
W=/tmp/c5 ; mkdir -p $W ; cp config.txt $W/
for topo in 8,16,32,16,8 16,16; do
for ch in 1 auto; do
  if [ $ch = auto ]; then unset LL_MMA_CHUNKS; else export LL_MMA_CHUNKS=$ch; fi
  ( cd $W && rm -rf data && timeout 600 /root/repo/lunar_lander_b200/host/bin/main_train_f --set NNConfig.hlayers=$topo --set NNConfig.epochs=300 > out.log 2> err.log; tail -1 err.log | sed "s/^/topo $topo chunks $ch: /" )
done; done
