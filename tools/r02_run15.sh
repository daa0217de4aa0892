#!/bin/bash
# one B200: render kernel v1 (convergent evaluation) against v2 (+ direct node load with band return, rcp-based skip
# quotients, zero-numerator divide shortcut): GPU tests on v2, then alternating bench lines
O=gpurun_out
ICB_LIB=$PWD/icon-ray-tracing_b200/libicon_b200_v2.so timeout 900 python -m pytest tests -m gpu -x -q > $O/r02n_pytest_v2.log 2>&1; tail -3 $O/r02n_pytest_v2.log
for v in v1 v2 v1 v2; do
  ICB_LIB=$PWD/icon-ray-tracing_b200/libicon_b200_$v.so timeout 300 python bench.py --cpu-seconds 0 --steps 20 > $O/r02n_bench_$v.json 2> $O/r02n_bench_$v.err
  python - <<P
import json
b=json.loads(open("$O/r02n_bench_$v.json").read().strip().splitlines()[-1])
print("$v", b["ms_per_step"], b["latency_ms"], b["config"]["samples_per_step"], b["config"]["slow_path_fraction"])
P
done
for w in "--sampler prism" "--integrator woodcock_grid"; do
for v in v1 v2; do
  ICB_LIB=$PWD/icon-ray-tracing_b200/libicon_b200_$v.so timeout 300 python bench.py --cpu-seconds 0 --steps 10 $w > $O/r02n_tmp.json 2> $O/r02n_tmp.err
  python - <<P
import json
b=json.loads(open("$O/r02n_tmp.json").read().strip().splitlines()[-1])
print("$w $v", b["ms_per_step"], b["latency_ms"])
P
done
done
