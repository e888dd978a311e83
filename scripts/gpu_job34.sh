#!/bin/bash
mkdir -p gpurun_out
python scripts/sanitize_small.py > gpurun_out/j34_plain.log 2>&1 && tail -3 gpurun_out/j34_plain.log
timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 9 python scripts/sanitize_small.py > gpurun_out/j34_memcheck.log 2>&1
echo "memcheck rc=$?"; grep -E "ERROR SUMMARY|Invalid|out of bounds" gpurun_out/j34_memcheck.log | head -10
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 9 python scripts/sanitize_small.py > gpurun_out/j34_racecheck.log 2>&1
echo "racecheck rc=$?"; grep -E "RACECHECK SUMMARY|hazard" gpurun_out/j34_racecheck.log | head -10
