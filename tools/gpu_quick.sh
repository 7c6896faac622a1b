#!/bin/bash
# Short gpurun call: GPU parity tests + solve profile (+ optional extra command).  Usage: bash tools/gpu_quick.sh tag ["extra command"]
TAG=${1:-quick}
O=gpurun_out/$TAG
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest_gpu.log
tail -15 $O/pytest_gpu.log
REPS=3 timeout 600 python tools/solve_profile.py 128 64 64 3 > $O/solve_128.log 2>&1; cat $O/solve_128.log
REPS=3 timeout 600 python tools/solve_profile.py 256 128 128 3 > $O/solve_256.log 2>&1; cat $O/solve_256.log
if [ -n "$2" ]; then bash -c "$2" > $O/extra.log 2>&1; tail -40 $O/extra.log; fi
