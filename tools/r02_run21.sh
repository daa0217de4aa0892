#!/bin/bash
# one B200, R2B10 x 90: L2 fetch granularity 32 / 64 (default) / 128 bytes
O=gpurun_out
export ICB_BUILD_TIMING=1
for gsz in 32 128; do
  ICB_L2_FETCH=$gsz timeout 400 python bench.py --workload C5 --cpu-seconds 0 --steps 8 --warmup 3 > $O/r02t_c5_l2fetch$gsz.json 2> $O/r02t_c5_l2fetch$gsz.err
  python - <<P
import json
b=json.loads(open("$O/r02t_c5_l2fetch$gsz.json").read().strip().splitlines()[-1]); print("l2fetch $gsz", round(b["ms_per_step"],2), round(b["latency_ms"],2), b["config"]["grid_build_s"])
P
done
grep 'buildGrid\|icb_set_grid' $O/r02t_c5_l2fetch128.err
for gsz in 32 128; do
  ICB_L2_FETCH=$gsz timeout 300 python bench.py --cpu-seconds 0 --steps 20 > $O/r02t_c3_l2fetch$gsz.json 2>/dev/null
  python - <<P
import json
b=json.loads(open("$O/r02t_c3_l2fetch$gsz.json").read().strip().splitlines()[-1]); print("C3 l2fetch $gsz", round(b["ms_per_step"],2), round(b["latency_ms"],2))
P
done
