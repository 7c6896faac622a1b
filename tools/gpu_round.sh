#!/bin/bash
# One gpurun call: GPU parity tests, bench (both arms), tile-shape sweep, ncu launch list + full capture of the K-apply kernel.
# Usage (from the repo root, on the GPU box):  bash tools/gpu_round.sh [tag]
TAG=${1:-run}
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/smoke.log
timeout 600 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" | tee -a $O/bench.err
timeout 300 python bench.py --impl reference --steps 20 --warmup 3 > $O/bench_ref.json 2> $O/bench_ref.err
SWEEP_CASES="-:-,4:-,0:-,3:-" timeout 600 python tools/march_sweep.py 128 64 64 > $O/sweep_128.log 2>&1
SWEEP_CASES="-:-,3:-,12:-,11:-" timeout 600 python tools/march_sweep.py 256 128 128 > $O/sweep_256.log 2>&1
REPS=3 timeout 600 python tools/solve_profile.py 128 64 64 3 > $O/solve_128.log 2>&1
REPS=3 timeout 600 python tools/solve_profile.py 256 128 128 3 > $O/solve_256.log 2>&1
# ncu: launch list of the bench command, then the K-apply kernel in full
BCMD="python bench.py --steps 20 --warmup 3 --no-cpu"
$BCMD > $O/plain_bench.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file $O/launches_bench.csv $BCMD > $O/ncu_bench.log 2>&1
ACMD="python bench.py --steps 6 --warmup 3 --no-cpu --no-solve"
$ACMD > $O/plain_apply.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:applyMarchKernel -s 4 -c 2 -o $O/apply_full -f $ACMD > $O/ncu_apply.log 2>&1
ls -la $O
