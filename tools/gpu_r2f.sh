#!/bin/bash
# 64x64-block factor kernel: parity + direct tests + refresh timing; full GPU test suite; bench
O=gpurun_out/r2f
mkdir -p $O
(timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_direct.py -x -q) > $O/parity_direct.log 2>&1; echo "rc=$?" >> $O/parity_direct.log
REPS=2 timeout 300 python tools/solve_profile.py 128 64 64 3 > $O/solve_128.log 2>&1
REPS=2 timeout 300 python tools/solve_profile.py 256 128 128 3 > $O/solve_256.log 2>&1
for b in 8 16 32 64 148; do echo "== VFEM_FACTOR_BLOCKS=$b"; VFEM_FACTOR_BLOCKS=$b REPS=1 timeout 300 python tools/solve_profile.py 128 64 64 3 | grep refresh; VFEM_FACTOR_BLOCKS=$b REPS=1 timeout 300 python tools/solve_profile.py 256 128 128 3 | grep refresh; done > $O/factor_blocks.log 2>&1
(time timeout 1500 python -m pytest tests -m gpu -x -q) > $O/all_gpu_tests.log 2>&1; echo "rc=$?" >> $O/all_gpu_tests.log
(time timeout 900 python bench.py --steps 20 --warmup 5) > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/bench.err
timeout 300 python __graft_entry__.py --smoke > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log
tail -n 5 $O/parity_direct.log $O/all_gpu_tests.log $O/smoke.log $O/bench.err; cat $O/solve_128.log $O/solve_256.log $O/factor_blocks.log
