#!/bin/bash
# host mirror (C++ trainer vs compiled reference) + nn_play side (stepped lander, cpp_nn_py, headless loop)
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests/test_gpu_host.py tests/test_gpu_play.py -q -m gpu ) > gpurun_out/j21_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/j21_pytest.log
tail -40 gpurun_out/j21_pytest.log
