#!/bin/bash
# One gpurun call: launch list + full captures of the three hot kernels on C3 (run from the repo root on the GPU box).
set -x
B="python bench.py --steps 2 --warmup 3 --cpu-seconds 0"
$B > gpurun_out/plain_march.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_march.csv $B > gpurun_out/ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:renderKernel -s 3 -c 1 -f -o gpurun_out/prof_march $B > gpurun_out/ncu_march.log 2>&1
$B --sampler prism > gpurun_out/plain_prism.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:renderKernel -s 3 -c 1 -f -o gpurun_out/prof_prism $B --sampler prism > gpurun_out/ncu_prism.log 2>&1
$B --integrator woodcock_grid > gpurun_out/plain_wgrid.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_wgrid.csv $B --integrator woodcock_grid > gpurun_out/ncu_launch2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:woodcockGridKernel -s 3 -c 1 -f -o gpurun_out/prof_wgrid $B --integrator woodcock_grid > gpurun_out/ncu_wgrid.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gridRangesKernel -c 1 -f -o gpurun_out/prof_ranges $B --integrator woodcock_grid > gpurun_out/ncu_ranges.log 2>&1
ls -la gpurun_out/*.ncu-rep
