#!/bin/bash
# Round-end evidence without the K-apply captures (kernel unchanged since r01r): smoke, bench (both arms), launch list of the
# bench command, ncu full capture of the stencil GS sweep kernel.  Usage: bash tools/gpu_final.sh [tag]
TAG=${1:-final}
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
timeout 300 python __graft_entry__.py --smoke > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/smoke.log
timeout 600 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" | tee -a $O/bench.err
timeout 300 python bench.py --impl reference --steps 20 --warmup 3 > $O/bench_ref.json 2> $O/bench_ref.err
BCMD="python bench.py --steps 20 --warmup 3 --no-cpu"
$BCMD > $O/plain_bench.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file $O/launches_bench.csv $BCMD > $O/ncu_bench.log 2>&1
SCMD="env NO_OPS=1 REPS=2 python tools/solve_profile.py 256 128 128 3"
$SCMD > $O/plain_solve.log 2>&1 &&
timeout 300 ncu --set full --clock-control none --import-source on -k regex:gsStencilSweepKernel -s 8 -c 2 -o $O/gs_sweep_full -f $SCMD > $O/ncu_sweep.log 2>&1
cat $O/bench.json; ls -la $O
