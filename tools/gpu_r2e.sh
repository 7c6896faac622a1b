#!/bin/bash
# plane-fused stencil GS (VFEM_STENCIL_PLANE=1): parity + timing; tile kernel with zero-filled box
O=gpurun_out/r2e
mkdir -p $O
(timeout 600 python -m pytest tests/test_gpu_parity.py -x -q) > $O/parity_tile0.log 2>&1; echo "rc=$?" >> $O/parity_tile0.log
(VFEM_STENCIL_PLANE=1 timeout 600 python -m pytest tests/test_gpu_parity.py -x -q) > $O/parity_sp.log 2>&1; echo "rc=$?" >> $O/parity_sp.log
if grep -q "rc=0" $O/parity_sp.log; then
(VFEM_STENCIL_PLANE=1 timeout 600 python -m pytest tests/test_gpu_slab.py -x -q) > $O/slab_sp.log 2>&1; echo "rc=$?" >> $O/slab_sp.log
for cy in 0 1 2 4 8; do echo "== stencil plane CY=$cy"; VFEM_GS_STPLANE_CY=$cy VFEM_STENCIL_PLANE=1 VFEM_GS_VARIANT=3 REPS=2 timeout 300 python tools/solve_profile.py 128 64 64 3 | grep "level\|solve 1\|FMG"; done > $O/gs_stplane.log 2>&1
for cy in 0 4 8; do echo "== 256 stencil plane CY=$cy"; VFEM_GS_STPLANE_CY=$cy VFEM_STENCIL_PLANE=1 VFEM_GS_VARIANT=3 REPS=2 timeout 300 python tools/solve_profile.py 256 128 128 3 | grep "level\|solve 1\|FMG"; done >> $O/gs_stplane.log 2>&1
fi
for v in 1 3; do echo "== tile variant $v (zero-filled box)"; VFEM_GS_VARIANT=$v REPS=2 timeout 300 python tools/solve_profile.py 128 64 64 3 | grep "level 0\|solve 1\|FMG";  VFEM_GS_VARIANT=$v REPS=1 timeout 300 python tools/solve_profile.py 256 128 128 3 | grep "level 0"; done >> $O/gs_stplane.log 2>&1
tail -n 6 $O/parity_tile0.log $O/parity_sp.log $O/slab_sp.log; cat $O/gs_stplane.log
