#!/bin/bash
# launch list + full capture of the default ray-gen kernel (woodcockGridKernel) in its state-machine form
set -x
B="python bench.py --integrator woodcock_grid --steps 2 --warmup 3 --cpu-seconds 0"
$B > gpurun_out/plain_wgrid3.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_wgrid3.csv $B > gpurun_out/ncu_launch_wgrid3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:woodcockGridKernel -s 3 -c 1 -f -o gpurun_out/prof_wgrid3 $B > gpurun_out/ncu_wgrid3.log 2>&1
ls -la gpurun_out/*.ncu-rep
