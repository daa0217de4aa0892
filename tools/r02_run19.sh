#!/bin/bash
# one B200: v2 against v4 (band prerequisites computed by all lanes together); GPU tests on v4; R2B10 x 90 capture
O=gpurun_out
for v in v2 v4 v2 v4; do
  ICB_LIB=$PWD/icon-ray-tracing_b200/libicon_b200_$v.so timeout 300 python bench.py --cpu-seconds 0 --steps 20 > $O/r02r_bench_$v.json 2> $O/r02r_bench_$v.err
  python - <<P
import json
b=json.loads(open("$O/r02r_bench_$v.json").read().strip().splitlines()[-1])
print("$v", b["ms_per_step"], b["latency_ms"], b["config"]["samples_per_step"])
P
done
cp icon-ray-tracing_b200/libicon_b200.so /tmp/keep.so; cp icon-ray-tracing_b200/libicon_b200_v4.so icon-ray-tracing_b200/libicon_b200.so
timeout 900 python -m pytest tests -m gpu -x -q > $O/r02r_pytest_v4.log 2>&1; tail -3 $O/r02r_pytest_v4.log
cp /tmp/keep.so icon-ray-tracing_b200/libicon_b200.so
# C5 (R2B10 x 90): why a third of C3's per-sample rate?  one full capture of the render kernel
B="python bench.py --workload C5 --steps 2 --warmup 3 --cpu-seconds 0 --in-flight 1"
timeout 600 ncu --set full --clock-control none -k regex:renderKernel -s 3 -c 1 -f -o $O/r02r_prof_march_c5 $B > $O/r02r_ncu_c5.log 2>&1
ls -la $O/r02r*.ncu-rep; tail -2 $O/r02r_ncu_c5.log
