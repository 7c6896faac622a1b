#!/bin/bash
# ncu captures; the .ncu-rep files are turned into text ON the box and deleted (gpurun_out is capped at 64 MiB)
O=gpurun_out/r2d
mkdir -p $O
timeout 120 tools/probes/fp64_probe > $O/fp64_probe.log 2>&1
export_rep() { # name
  ncu -i $O/$1.ncu-rep --page raw --csv > $O/$1_raw.csv 2>/dev/null
  ncu -i $O/$1.ncu-rep --page details > $O/$1_details.txt 2>/dev/null
  ncu -i $O/$1.ncu-rep --page source --csv --print-source sass > $O/$1_source_sass.csv 2>/dev/null
  rm -f $O/$1.ncu-rep
  gzip -f $O/$1_source_sass.csv
}
SCMD="python tools/solve_profile.py 128 64 64 3"
NO_OPS=1 REPS=1 $SCMD > $O/plain_solve.log 2>&1 &&
NO_OPS=1 REPS=1 timeout 900 ncu --set full --clock-control none -k regex:gsFineTileKernel -s 16 -c 1 -o $O/gs_tile_full -f $SCMD > $O/ncu_gs.log 2>&1
export_rep gs_tile_full
ACMD="python bench.py --steps 6 --warmup 3 --no-cpu --no-solve --no-extra"
$ACMD > $O/plain_apply.log 2>&1 &&
timeout 900 ncu --set full --clock-control none -k regex:applyMarchTmaKernel -s 40 -c 1 -o $O/apply_full -f $ACMD > $O/ncu_apply.log 2>&1
export_rep apply_full
NO_OPS=1 REPS=2 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $O/launches_solve.csv $SCMD > $O/ncu_launches.log 2>&1
NO_OPS=1 REPS=1 timeout 900 ncu --set full --clock-control none -k regex:blockTriFactorKernel -c 1 -o $O/factor_full -f $SCMD > $O/ncu_factor.log 2>&1
export_rep factor_full
for v in 1 2 3 4; do echo "== VFEM_GS_VARIANT=$v"; VFEM_GS_VARIANT=$v REPS=2 timeout 300 python tools/solve_profile.py 128 64 64 3 | grep "level 0\|solve 1\|FMG"; VFEM_GS_VARIANT=$v REPS=1 timeout 300 python tools/solve_profile.py 256 128 128 3 | grep "level 0"; done > $O/gs_variants.log 2>&1
grep MIX $O/fp64_probe.log; cat $O/gs_variants.log; ls -la $O; du -sh $O
