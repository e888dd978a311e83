#!/bin/bash
# small populations: neurons of a slot dealt to 4 lanes by default: full GPU suite + configs[0] bench line
mkdir -p gpurun_out
O=gpurun_out
( time timeout 1800 python -m pytest tests -q -m gpu ) > $O/j45_pytest.log 2>&1
echo "pytest rc=$?" >> $O/j45_pytest.log
tail -6 $O/j45_pytest.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3
timeout 600 python bench.py --workload c1 > $O/j45_bench_c1.json 2> $O/j45_bench_c1.err
cut -c1-220 $O/j45_bench_c1.json
timeout 600 python bench.py --workload c1 --precision 32 --exact --no-cpu-baseline --e2e-steps 0 2>/dev/null | cut -c1-220
W=/tmp/c1run ; mkdir -p $W ; cp config.txt $W/
( cd $W && ( time /root/repo/lunar_lander_b200/host/bin/main_train_d --set NNConfig.epochs=200 > out.log 2> err.log ) 2> time.log ; tail -3 out.log; tail -3 time.log )
