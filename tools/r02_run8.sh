#!/bin/bash
# occupancy variants of the render kernel (register caps by __maxnreg__): 3 x 192 threads at 112 registers, 2 x 320 at 104
O=gpurun_out
for v in v192 v320; do
  ICB_LIB=$PWD/icon-ray-tracing_b200/libicon_b200_$v.so timeout 600 python bench.py --cpu-seconds 0 --steps 20 > $O/r02h_bench_$v.json 2> $O/r02h_bench_$v.err; tail -c 300 $O/r02h_bench_$v.json; tail -2 $O/r02h_bench_$v.err
done
timeout 600 python bench.py --cpu-seconds 0 --steps 20 > $O/r02h_bench_base.json 2> /dev/null
