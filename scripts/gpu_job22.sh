#!/bin/bash
mkdir -p gpurun_out /tmp/dd && cd /tmp/dd
python - <<'P'
import sys; sys.path.insert(0, "/root/repo/tests")
from test_gpu_host import write_config, SMALL
ov = dict(SMALL); ov.update({"NNConfig.hlayers": "8,16,32,16,8", "NNConfig.epochs": 6, "NNConfig.population_size": 96})
for d in ("a", "b", "c"): write_config("/tmp/dd/" + d, **ov)
P
(cd a && /root/repo/lunar_lander_b200/host/bin/main_train_d --set GpuCfg.select=device > log.txt 2>err.txt)
(cd b && /root/repo/lunar_lander_b200/host/bin/main_train_d --set GpuCfg.select=host > log.txt 2>err.txt)
(cd c && /root/repo/oracle/_ref/main_train_d > log.txt 2>err.txt)
diff a/log.txt b/log.txt; echo "--- dev vs ref"; diff a/log.txt c/log.txt; echo "--- host vs ref"; diff b/log.txt c/log.txt
cat a/log.txt > /root/repo/gpurun_out/j22_a.log
