#!/bin/bash
# one B200: ICB_REFILL_MIN (idle lanes that trigger a warp's refill) with the round-end kernel
for r in 8 4 12 16 8; do
  ICB_REFILL_MIN=$r timeout 40 python bench.py --cpu-seconds 0 --steps 20 2>/dev/null | python -c "import sys,json; b=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('refill_min $r', round(b['ms_per_step'],3), round(b['latency_ms'],3))"
done
