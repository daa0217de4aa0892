#!/bin/bash
# one B200: the default ray-gen with 32-bit macrocell indices (v5) against the shipped build, GPU tests of the Woodcock kernels on v5
O=gpurun_out
for v in base v5 base v5; do
  lib=$PWD/icon-ray-tracing_b200/libicon_b200_v5.so; [ $v = base ] && lib=$PWD/icon-ray-tracing_b200/libicon_b200.so
  ICB_LIB=$lib timeout 300 python bench.py --integrator woodcock_grid --cpu-seconds 0 --steps 20 > $O/r02u_wgrid_$v.json 2>/dev/null
  python - <<P
import json
b=json.loads(open("$O/r02u_wgrid_$v.json").read().strip().splitlines()[-1]); print("wgrid $v", round(b["ms_per_step"],3), round(b["latency_ms"],3))
P
done
cp icon-ray-tracing_b200/libicon_b200_v5.so icon-ray-tracing_b200/libicon_b200.so
timeout 600 python -m pytest tests/test_woodcock_grid.py tests/test_woodcock_sphere.py tests/test_gpu_parity.py -m gpu -x -q > $O/r02u_pytest_v5.log 2>&1; tail -2 $O/r02u_pytest_v5.log
