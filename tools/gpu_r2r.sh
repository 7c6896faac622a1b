#!/bin/bash
# ncu --set full of the level-1 stencil sweep (largest share of a solve) and of the new Galerkin kernel; text exports only
O=gpurun_out/r2r
mkdir -p $O
export_rep() { ncu -i $O/$1.ncu-rep --page details > $O/$1_details.txt 2>/dev/null; ncu -i $O/$1.ncu-rep --page raw --csv > $O/$1_raw.csv 2>/dev/null; rm -f $O/$1.ncu-rep; }
SCMD="python tools/solve_profile.py 128 64 64 3"
NO_OPS=1 REPS=1 $SCMD > $O/plain_solve.log 2>&1 &&
NO_OPS=1 REPS=1 timeout 600 ncu --set full --clock-control none -k regex:gsStencilSweepKernel -s 2 -c 1 -o $O/stencil_sweep_full -f $SCMD > $O/ncu_sweep.log 2>&1
export_rep stencil_sweep_full
NO_OPS=1 REPS=1 timeout 600 ncu --set full --clock-control none -k regex:galerkinStencilCtaKernel -c 1 -o $O/galerkin_full -f $SCMD > $O/ncu_gal.log 2>&1
export_rep galerkin_full
ls -la $O; grep -i "duration\|Registers Per\|Achieved Occupancy\|DRAM Throughput\|Issue Slots Busy\|Memory Throughput" $O/stencil_sweep_full_details.txt | head -12; grep -i "duration\|DRAM Throughput\|Issue Slots Busy" $O/galerkin_full_details.txt | head -6
