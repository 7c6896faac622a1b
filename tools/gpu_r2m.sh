#!/bin/bash
# Galerkin kernel forms (0 thread per block, 1 warp per block, 2 CTA per coarse node): parity, refresh timing
O=gpurun_out/r2m
mkdir -p $O
(timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_direct.py tests/test_gpu_slab.py tests/test_gpu_baseline.py -x -q) > $O/tests.log 2>&1; echo "rc=$?" >> $O/tests.log
for v in 2 1 0; do
  export VFEM_GALERKIN_FORM=$v
  echo "== form $v"; REPS=2 timeout 300 python tools/solve_profile.py 128 64 64 3 | grep "solve 1\|coarse"; REPS=2 timeout 300 python tools/solve_profile.py 256 128 128 3 | grep "solve 1\|coarse"
done > $O/refresh.log 2>&1
tail -n 5 $O/tests.log; cat $O/refresh.log
