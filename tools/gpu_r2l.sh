#!/bin/bash
# e2e pipeline depth sweep, final bench record, ncu launch list of the bench command
O=gpurun_out/r2l
mkdir -p $O
timeout 300 python tools/e2e_probe.py > $O/e2e_probe.log 2>&1
(time timeout 600 python bench.py) > $O/bench.json 2> $O/bench.err; echo "rc=$?" >> $O/bench.err
(time timeout 300 python bench.py --impl reference) > $O/bench_ref.json 2> $O/bench_ref.err; echo "rc=$?" >> $O/bench_ref.err
CMD="python bench.py --steps 20 --warmup 5 --no-cpu --no-extra"
VFEM_BENCH_PREHEAT_S=0 $CMD > $O/plain_short.json 2> $O/plain_short.err &&
VFEM_BENCH_PREHEAT_S=0 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/launches_bench.csv $CMD > $O/ncu_bench.log 2>&1
cat $O/e2e_probe.log; grep "rc=" $O/*.err; head -c 600 $O/bench.json; echo; wc -l $O/launches_bench.csv
