#!/bin/bash
# ncu captures (kept small: gpurun_out is capped at 64 MiB): ring-fed march kernel, tile GS kernel, launch list of one solve
O=gpurun_out/r2d
mkdir -p $O
timeout 120 tools/probes/fp64_probe > $O/fp64_probe.log 2>&1
SCMD="python tools/solve_profile.py 128 64 64 3"
NO_OPS=1 REPS=1 $SCMD > $O/plain_solve.log 2>&1 &&
NO_OPS=1 REPS=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:gsFineTileKernel -s 16 -c 1 -o $O/gs_tile_full -f $SCMD > $O/ncu_gs.log 2>&1
ACMD="python bench.py --steps 6 --warmup 3 --no-cpu --no-solve --no-extra"
$ACMD > $O/plain_apply.log 2>&1 &&
timeout 900 ncu --set full --clock-control none -k regex:applyMarchTmaKernel -s 40 -c 1 -o $O/apply_full -f $ACMD > $O/ncu_apply.log 2>&1
NO_OPS=1 REPS=2 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $O/launches_solve.csv $SCMD > $O/ncu_launches.log 2>&1
NO_OPS=1 REPS=1 timeout 900 ncu --section SpeedOfLight --section WarpStateStats --section Occupancy --section LaunchStats --clock-control none -k regex:blockTriFactorKernel -c 1 -o $O/factor_sol -f $SCMD > $O/ncu_factor.log 2>&1
cat $O/fp64_probe.log; ls -la $O; du -sh $O
