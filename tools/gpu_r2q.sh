#!/bin/bash
# grid chain solve (P > 256): CTA size sweep; parity of the paths that use it
O=gpurun_out/r2q
mkdir -p $O
(timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_direct.py -x -q) > $O/tests.log 2>&1; echo "rc=$?" >> $O/tests.log
for t in default 256; do
  [ $t != default ] && export VFEM_CHAIN_GRID_THREADS=$t
  echo "== threads $t"; REPS=2 timeout 300 python tools/solve_profile.py 256 128 128 3 | grep "solve 1\|coarse"
done > $O/chain.log 2>&1
tail -n 4 $O/tests.log; cat $O/chain.log
