#!/bin/bash
# one B200: GPU tests on the new builder + kernel, the render kernel before/after the convergent post-fetch evaluation,
# then BASELINE configs[4] and configs[3] on one GPU
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $O/r02j_pytest.log 2>&1; tail -3 $O/r02j_pytest.log
for v in base v1 base v1; do
  ICB_LIB=$PWD/icon-ray-tracing_b200/libicon_b200_$v.so timeout 300 python bench.py --cpu-seconds 0 --steps 20 > $O/r02j_bench_$v.json 2> $O/r02j_bench_$v.err
  python - <<P
import json
b=json.loads(open("$O/r02j_bench_$v.json").read().strip().splitlines()[-1])
print("$v", b["ms_per_step"], b["latency_ms"], b["config"]["samples_per_step"])
P
done
bash tools/run_configs.sh 1
