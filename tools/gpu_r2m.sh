#!/bin/bash
# Galerkin kernel (CTA per coarse node; VFEM_GALERKIN_THREAD=1: round-1 thread form): parity, refresh timing
O=gpurun_out/r2m
mkdir -p $O
(timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_direct.py tests/test_gpu_slab.py tests/test_gpu_baseline.py -x -q) > $O/tests.log 2>&1; echo "rc=$?" >> $O/tests.log
(REPS=2 timeout 300 python tools/solve_profile.py 128 64 64 3 | grep "solve 1\|coarse"; REPS=2 timeout 300 python tools/solve_profile.py 256 128 128 3 | grep "solve 1\|coarse") > $O/refresh.log 2>&1
tail -n 3 $O/tests.log; cat $O/refresh.log
