#!/bin/bash
# Parity suite + solve profiles at both grids.  Usage: bash tools/gpu_sweep.sh tag
TAG=${1:-sweep}
O=gpurun_out/$TAG
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest_gpu.log
tail -5 $O/pytest_gpu.log
for g in "128 64 64" "256 128 128"; do
  echo "# grid $g" | tee -a $O/solve.log
  REPS=3 timeout 600 python tools/solve_profile.py $g 3 2>&1 | tee -a $O/solve.log
done
