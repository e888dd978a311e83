#!/bin/bash
# BASELINE configs[3] at full size: 2^20 members of 32_128_32 x 50 layouts over 8 GPUs
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 8 --steps 3 --warmup 3 > gpurun_out/j42_bench_c4_8gpu.json 2> gpurun_out/j42_bench_c4_8gpu.err
cut -c1-400 gpurun_out/j42_bench_c4_8gpu.json; tail -3 gpurun_out/j42_bench_c4_8gpu.err
