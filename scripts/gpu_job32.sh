#!/bin/bash
# the sharded Python trainer: 2 ranks must print the log of 1 rank (the GA does not depend on the GPU count)
mkdir -p gpurun_out /tmp/t1 /tmp/t2
O=/root/repo/gpurun_out
for d in /tmp/t1 /tmp/t2; do cp config.txt $d/; done
ARGS="--float --set NNConfig.hlayers=16,64,32 --set NNConfig.population_size=4096 --set NNConfig.epochs=12 --set NNTrainingCfg.layout_nb=20 --set NNConfig.save_interval=6"
( cd /tmp/t1 && PYTHONPATH=/root/repo python -m lunar_lander_b200.trainer $ARGS > $O/j32_trainer_1rank.log 2> $O/j32_trainer_1rank.err )
( cd /tmp/t2 && PYTHONPATH=/root/repo python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 -m lunar_lander_b200.trainer $ARGS > $O/j32_trainer_2rank.log 2> $O/j32_trainer_2rank.err )
tail -7 $O/j32_trainer_2rank.log; tail -3 $O/j32_trainer_2rank.err
diff <(sed -E 's/\(Took: [0-9:]+, Total: [0-9:]+\)//' $O/j32_trainer_1rank.log) <(sed -E 's/\(Took: [0-9:]+, Total: [0-9:]+\)//' $O/j32_trainer_2rank.log) && echo "LOGS IDENTICAL (1 rank vs 2 ranks)"
for f in /tmp/t1/data/*; do cmp $f /tmp/t2/data/$(basename $f) && echo "same $(basename $f)"; done
timeout 600 python -m pytest tests/test_gpu_host.py -q -m gpu 2>&1 | tail -3
