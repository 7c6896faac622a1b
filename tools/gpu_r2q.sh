#!/bin/bash
# chain solves (cluster kernel P <= 256 / grid kernel): parity of the paths that use them, timings
O=gpurun_out/r2q
mkdir -p $O
(timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_direct.py tests/test_gpu_baseline.py -x -q) > $O/tests.log 2>&1; echo "rc=$?" >> $O/tests.log
(REPS=2 timeout 300 python tools/solve_profile.py 128 64 64 3 | grep "solve 1\|coarse"; REPS=2 timeout 300 python tools/solve_profile.py 256 128 128 3 | grep "solve 1\|coarse") > $O/chain.log 2>&1
tail -n 4 $O/tests.log; cat $O/chain.log
