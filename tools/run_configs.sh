#!/bin/bash
# BASELINE configs[3] (C4, time series) and configs[4] (C5, R2B10 x 90) on N GPUs of one box
N=${1:-8}
O=gpurun_out
bash tools/run_c5_multi.sh $N 8
if [ "$N" = "1" ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29542"; fi
timeout 600 $L bench.py --gpus $N --workload C4 --cpu-seconds 0 --steps 24 --warmup 3 --timesteps 24 > $O/c4_n${N}.json 2> $O/c4_n${N}.err
tail -c 1500 $O/c4_n${N}.json; grep -v Warning $O/c4_n${N}.err | tail -5
