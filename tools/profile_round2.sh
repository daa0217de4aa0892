#!/bin/bash
# launch list of the default bench command + full capture of the prism kernel (current state)
set -x
B="python bench.py --steps 2 --warmup 3 --cpu-seconds 0"
$B > gpurun_out/plain_default.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_default.csv $B > gpurun_out/ncu_launch_default.log 2>&1
$B --sampler prism --in-flight 1 > gpurun_out/plain_prism2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:renderKernel -s 3 -c 1 -f -o gpurun_out/prof_prism2 $B --sampler prism --in-flight 1 > gpurun_out/ncu_prism2.log 2>&1
ls -la gpurun_out/*.ncu-rep
