#!/bin/bash
# four B200s: the default bench line at N = 4 and N = 2 with the round-end build
O=gpurun_out
for N in 4 2; do
  L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2954$N"
  timeout 300 $L bench.py --gpus $N --steps 20 --warmup 5 > $O/r02v_bench_n$N.json 2> $O/r02v_bench_n$N.err
  python - <<P
import json
b=json.loads(open("$O/r02v_bench_n$N.json").read().strip().splitlines()[-1]); print("N=$N", round(b["ms_per_step"],3), round(b["latency_ms"],3), b["value"]/1e9, b["config"]["frame_identical_to_single_gpu"], b["e2e"]["ms_per_step"])
P
done
