#!/bin/bash
# round 2, first call: parity tests with the tile GS kernel + hand-written factorisation, direct-solve goldens, BASELINE-config tests,
# solve profiles (tile vs gather GS), the rewritten bench
O=gpurun_out/r2a
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/smi.txt; nproc >> $O/smi.txt
(time timeout 600 python -m pytest tests/test_gpu_parity.py -x -q) > $O/parity_tests.log 2>&1; echo "rc=$?" >> $O/parity_tests.log
(time timeout 600 python -m pytest tests/test_gpu_direct.py tests/test_boundary_shims.py -m gpu -q) > $O/direct_tests.log 2>&1; echo "rc=$?" >> $O/direct_tests.log
(time timeout 900 python -m pytest tests/test_gpu_march.py tests/test_gpu_topopt.py tests/test_gpu_torch_loop.py -x -q) > $O/more_tests.log 2>&1; echo "rc=$?" >> $O/more_tests.log
REPS=3 timeout 600 python tools/solve_profile.py 128 64 64 3 > $O/solve_128_tile.log 2>&1
VFEM_GS_VARIANT=0 REPS=3 timeout 600 python tools/solve_profile.py 128 64 64 3 > $O/solve_128_gather.log 2>&1
REPS=2 timeout 600 python tools/solve_profile.py 256 128 128 3 > $O/solve_256_tile.log 2>&1
(time timeout 1500 python -m pytest tests/test_gpu_baseline.py -q -s) > $O/baseline_tests.log 2>&1; echo "rc=$?" >> $O/baseline_tests.log
(time timeout 900 python bench.py --steps 20 --warmup 5) > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/bench.err
(time timeout 300 python bench.py --steps 20 --warmup 5 --no-extra --no-cpu --no-solve) > $O/bench2.json 2>> $O/bench.err
(time timeout 300 python bench.py --impl reference --steps 20 --warmup 5) > $O/bench_ref.json 2>> $O/bench.err
(time timeout 900 python -m pytest tests/test_gpu_slab.py -x -q) > $O/slab_tests.log 2>&1; echo "rc=$?" >> $O/slab_tests.log
tail -4 $O/parity_tests.log $O/direct_tests.log $O/more_tests.log $O/baseline_tests.log $O/slab_tests.log; cat $O/solve_128_tile.log $O/solve_128_gather.log $O/solve_256_tile.log
