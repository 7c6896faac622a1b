#!/bin/bash
# Fine-level GS: split kernel (VFEM_GS_VARIANT=8) vs one-thread-per-node (4).  Usage: bash tools/gpu_gs_split.sh tag
TAG=${1:-gs_split}
O=gpurun_out/$TAG
mkdir -p $O
VFEM_GS_VARIANT=${GSV:-8} timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_v8.log 2>&1; echo "pytest(v8) rc=$?" | tee -a $O/pytest_gpu_v8.log
tail -5 $O/pytest_gpu_v8.log
for v in 4 ${GSV:-8}; do
  for g in "128 64 64" "256 128 128"; do
    echo "# variant $v grid $g" | tee -a $O/solve.log
    VFEM_GS_VARIANT=$v REPS=3 timeout 600 python tools/solve_profile.py $g 3 2>&1 | tee -a $O/solve.log
  done
done
