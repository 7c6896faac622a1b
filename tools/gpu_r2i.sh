#!/bin/bash
# the allocation-free entry point: alone in a fresh process, then after the whole GPU suite
O=gpurun_out/r2i
mkdir -p $O
(timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q) > $O/no_alloc_alone.log 2>&1; echo "rc=$?" >> $O/no_alloc_alone.log
(time timeout 1500 python -m pytest tests -m gpu -x -q) > $O/all_gpu_tests.log 2>&1; echo "rc=$?" >> $O/all_gpu_tests.log
(timeout 300 python -c "import __graft_entry__ as g; g.smoke()") > $O/smoke.log 2>&1; echo "rc=$?" >> $O/smoke.log
tail -n 4 $O/no_alloc_alone.log; tail -n 8 $O/all_gpu_tests.log; tail -n 3 $O/smoke.log
