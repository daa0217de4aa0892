#!/bin/bash
# round-end check on one B200: GPU tests, smoke(), the default bench line, the reference arm, the other kernels' lines,
# then (after those exited 0) the ncu launch list of the default command and a full capture of the default ray-gen kernel
set -x
O=gpurun_out
python -m pytest tests -m gpu -x -q > $O/final_pytest.log 2>&1; tail -2 $O/final_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/final_smoke.log 2>&1; tail -1 $O/final_smoke.log
python bench.py > $O/final_n1.json 2> $O/final_n1.err; tail -c 300 $O/final_n1.json
python bench.py --impl reference --steps 3 --warmup 1 > $O/final_ref.json 2> $O/final_ref.err; tail -c 300 $O/final_ref.json
python bench.py --sampler prism --cpu-seconds 0 > $O/final_prism.json 2>/dev/null
python bench.py --integrator woodcock_grid --cpu-seconds 0 > $O/final_wgrid.json 2>/dev/null
python bench.py --in-flight 1 --cpu-seconds 0 > $O/final_inflight1.json 2>/dev/null
B="python bench.py --steps 2 --warmup 3 --cpu-seconds 0"
$B > $O/plain_default3.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_default3.csv $B > $O/ncu_launch_default3.log 2>&1
W="python bench.py --integrator woodcock_grid --steps 2 --warmup 3 --cpu-seconds 0"
$W > $O/plain_wgrid4.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_wgrid4.csv $W > $O/ncu_launch_wgrid4.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:woodcockGridKernel -s 3 -c 1 -f -o $O/prof_wgrid4 $W > $O/ncu_wgrid4.log 2>&1
ls -la $O/*.ncu-rep
