#!/bin/bash
# round 2, one B200: new tests (sphere ray-gen, CLI), bench lines of every kernel, launch lists and one full ncu capture per kernel
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $O/r02e_pytest.log 2>&1; tail -4 $O/r02e_pytest.log
B="python bench.py --cpu-seconds 0"
timeout 600 python bench.py --steps 20 --warmup 3 > $O/r02e_bench_n1.json 2> $O/r02e_bench_n1.err; tail -c 400 $O/r02e_bench_n1.json
ICB_NO_FAST_PLANES=1 timeout 600 $B --steps 20 > $O/r02e_bench_n1_exactplanes.json 2> /dev/null; tail -c 300 $O/r02e_bench_n1_exactplanes.json
timeout 600 $B --steps 20 --in-flight 8 > $O/r02e_bench_n1_if8.json 2> /dev/null; tail -c 300 $O/r02e_bench_n1_if8.json
for v in "--sampler prism:prism" "--integrator woodcock_grid:wgrid" "--integrator woodcock_sphere:wsphere" "--integrator woodcock --steps 4:wae"; do
  a="${v%%:*}"; n="${v##*:}"
  timeout 600 $B $a > $O/r02e_bench_$n.json 2> $O/r02e_bench_$n.err; tail -c 300 $O/r02e_bench_$n.json; tail -2 $O/r02e_bench_$n.err
done
# launch list of the default command, then one full capture per kernel (frames one at a time: the launch alone on the device)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02e_launches_default.csv python bench.py --steps 2 --warmup 3 --cpu-seconds 0 --min-region 0 > $O/r02e_ncu_launch.log 2>&1
P="--steps 2 --warmup 3 --cpu-seconds 0 --in-flight 1 --min-region 0"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:renderKernel -s 4 -c 1 -f -o $O/r02e_prof_march python bench.py $P > $O/r02e_ncu_march.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:renderKernel -s 4 -c 1 -f -o $O/r02e_prof_prism python bench.py $P --sampler prism > $O/r02e_ncu_prism.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:woodcockGridKernel -s 3 -c 1 -f -o $O/r02e_prof_wgrid python bench.py $P --integrator woodcock_grid > $O/r02e_ncu_wgrid.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:woodcockSphereKernel -s 3 -c 1 -f -o $O/r02e_prof_wsphere python bench.py $P --integrator woodcock_sphere > $O/r02e_ncu_wsphere.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:woodcockKernel -s 3 -c 1 -f -o $O/r02e_prof_wae python bench.py --steps 1 --warmup 3 --cpu-seconds 0 --in-flight 1 --min-region 0 --integrator woodcock > $O/r02e_ncu_wae.log 2>&1
ls -la $O/*.ncu-rep
