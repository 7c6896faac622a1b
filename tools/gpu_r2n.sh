#!/bin/bash
# full GPU suite with the final kernels (+ the 50-iteration regression trajectory written once), bench, smoke
O=gpurun_out/r2n
mkdir -p $O
(time VFEM_WRITE_GOLDEN=$O/config2_oc50_gpu.json timeout 1700 python -m pytest tests -m gpu -x -q --durations=8) > $O/all_gpu_tests.log 2>&1; echo "rc=$?" >> $O/all_gpu_tests.log
(timeout 300 python -c "import __graft_entry__ as g; g.smoke()") > $O/smoke.log 2>&1; echo "rc=$?" >> $O/smoke.log
(time timeout 600 python bench.py) > $O/bench.json 2> $O/bench.err; echo "rc=$?" >> $O/bench.err
tail -n 20 $O/all_gpu_tests.log; tail -n 2 $O/smoke.log; grep rc= $O/bench.err; head -c 400 $O/bench.json
