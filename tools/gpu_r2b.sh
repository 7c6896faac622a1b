#!/bin/bash
# round 2, call b: ring-fed march kernel, rewritten factor kernel, tile GS: tests, timings, ncu captures
O=gpurun_out/r2b
mkdir -p $O
(time timeout 600 python -m pytest tests/test_gpu_march.py tests/test_gpu_parity.py -x -q) > $O/march_parity_tests.log 2>&1; echo "rc=$?" >> $O/march_parity_tests.log
(time timeout 600 python -m pytest tests/test_gpu_direct.py tests/test_boundary_shims.py tests/test_gpu_topopt.py -m gpu -q) > $O/direct_tests.log 2>&1; echo "rc=$?" >> $O/direct_tests.log
REPS=3 timeout 600 python tools/solve_profile.py 128 64 64 3 > $O/solve_128.log 2>&1
REPS=2 timeout 600 python tools/solve_profile.py 256 128 128 3 > $O/solve_256.log 2>&1
for v in 0 4; do
  VFEM_APPLY_VARIANT=$v timeout 600 python bench.py --steps 50 --warmup 5 --no-cpu --no-solve > $O/bench_variant$v.json 2>> $O/bench.err
done
SWEEP_CASES="-:-,4:-,8:-,5:-,0:-,3:-,11:-" timeout 600 python tools/march_sweep.py 128 64 64 > $O/sweep_128.log 2>&1
SWEEP_CASES="-:-,3:-,12:-,11:-,0:-,4:-" timeout 600 python tools/march_sweep.py 256 128 128 > $O/sweep_256.log 2>&1
(time timeout 900 python -m pytest tests/test_gpu_baseline.py tests/test_gpu_slab.py tests/test_gpu_torch_loop.py -q) > $O/baseline_slab_tests.log 2>&1; echo "rc=$?" >> $O/baseline_slab_tests.log
# ncu: tile GS kernel, factor kernel, march kernel (full sets), then the launch list of one solve
SCMD="python tools/solve_profile.py 128 64 64 3"
NO_OPS=1 REPS=1 $SCMD > $O/plain_solve.log 2>&1 &&
NO_OPS=1 REPS=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:gsFineTileKernel -s 16 -c 2 -o $O/gs_tile_full -f $SCMD > $O/ncu_gs.log 2>&1
NO_OPS=1 REPS=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:blockTriFactorKernel -c 1 -o $O/factor_full -f $SCMD > $O/ncu_factor.log 2>&1
ACMD="python bench.py --steps 6 --warmup 3 --no-cpu --no-solve --no-extra"
$ACMD > $O/plain_apply.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:applyMarchTmaKernel -s 40 -c 2 -o $O/apply_full -f $ACMD > $O/ncu_apply.log 2>&1
NO_OPS=1 REPS=2 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $O/launches_solve.csv $SCMD > $O/ncu_launches.log 2>&1
tail -n 4 $O/march_parity_tests.log $O/direct_tests.log $O/baseline_slab_tests.log; cat $O/solve_128.log $O/solve_256.log $O/sweep_128.log $O/sweep_256.log; cat $O/bench_variant*.json | cut -c1-400
ls -la $O
