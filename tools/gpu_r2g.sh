#!/bin/bash
# full GPU suite + solve timings after the factor-kernel revert
O=gpurun_out/r2g
mkdir -p $O
(time timeout 1500 python -m pytest tests -m gpu -x -q) > $O/all_gpu_tests.log 2>&1; echo "rc=$?" >> $O/all_gpu_tests.log
REPS=2 timeout 300 python tools/solve_profile.py 128 64 64 3 > $O/solve_128.log 2>&1
REPS=2 timeout 300 python tools/solve_profile.py 256 128 128 3 > $O/solve_256.log 2>&1
tail -n 8 $O/all_gpu_tests.log; grep "solve 1\|coarse" $O/solve_128.log $O/solve_256.log
