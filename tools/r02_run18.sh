#!/bin/bash
# eight B200s, C3: frames in flight x CTA share x smallest per-frame grid, then the default bench line at N = 8
O=gpurun_out
L="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29544"
timeout 300 $L tools/sweep_scale.py --configs "8,4,0,384,16,10;12,4,0,384,16,10;16,4,0,384,16,10;4,2,0,384,16,10" > $O/r02q_sweep_n8_min74.jsonl 2> $O/r02q_sweep_a.err
grep ms_per_frame $O/r02q_sweep_n8_min74.jsonl | cut -c1-160
ICB_MIN_BLOCKS=37 timeout 300 $L tools/sweep_scale.py --configs "8,8,0,384,16,10;12,8,0,384,16,10;16,8,0,384,16,10;16,16,0,384,16,10" > $O/r02q_sweep_n8_min37.jsonl 2> $O/r02q_sweep_b.err
grep ms_per_frame $O/r02q_sweep_n8_min37.jsonl | cut -c1-160
timeout 300 $L bench.py --gpus 8 --steps 20 --warmup 5 > $O/r02q_bench_n8.json 2> $O/r02q_bench_n8.err
tail -c 1200 $O/r02q_bench_n8.json; grep -v Warn $O/r02q_bench_n8.err | tail -3
