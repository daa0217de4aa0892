#!/bin/bash
# one B200: v2 against v3 (sRGB packing out of line), then launch list + full capture of the shipped kernel
O=gpurun_out
for v in v2 v3 v2 v3; do
  ICB_LIB=$PWD/icon-ray-tracing_b200/libicon_b200_$v.so timeout 300 python bench.py --cpu-seconds 0 --steps 20 > $O/r02p_bench_$v.json 2> $O/r02p_bench_$v.err
  python - <<P
import json
b=json.loads(open("$O/r02p_bench_$v.json").read().strip().splitlines()[-1])
print("$v", b["ms_per_step"], b["latency_ms"], b["config"]["samples_per_step"])
P
done
B="python bench.py --steps 2 --warmup 3 --cpu-seconds 0 --in-flight 1"
$B > $O/r02p_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02p_launches_march_c3.csv $B > $O/r02p_ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:renderKernel -s 3 -c 1 -f -o $O/r02p_prof_march $B > $O/r02p_ncu_full.log 2>&1
ls -la $O/r02p*.ncu-rep
