#!/bin/bash
# weak scaling of configs[3] at 2 and 4 GPUs on the single-accumulator kernel
mkdir -p gpurun_out
N=$1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2954$N bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/j46_bench_c4_${N}gpu.json 2> gpurun_out/j46_bench_c4_${N}gpu.err
cut -c1-300 gpurun_out/j46_bench_c4_${N}gpu.json; tail -2 gpurun_out/j46_bench_c4_${N}gpu.err
