#!/bin/bash
# round 2, eight B200: what limits the sort-first frame rate at N=8 -- frames in flight beyond 8, kernel grid floor, comm stream
# priority, size of the dealt blocks; then the time series (C4) and the bench line at N=8
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 8"
S="tools/sweep_scale.py --frames 20"
timeout 300 $TR --master-port 29517 $S --configs "8,8,0,384,16,10;12,8,0,384,16,10;16,8,0,384,16,10;16,16,0,384,16,10;12,4,0,384,16,10" > $O/r02f_sweep_n8_slots.jsonl 2> $O/r02f_sweep_a.err; cat $O/r02f_sweep_n8_slots.jsonl
ICB_MIN_BLOCKS=37 timeout 300 $TR --master-port 29518 $S --configs "8,8,0,384,16,10;16,8,0,384,16,10;16,16,0,384,16,10" > $O/r02f_sweep_n8_minblocks37.jsonl 2> $O/r02f_sweep_b.err; cat $O/r02f_sweep_n8_minblocks37.jsonl
ICB_DEAL_BLOCK=64 timeout 300 $TR --master-port 29519 $S --configs "8,8,0,384,16,10;16,8,0,384,16,10" > $O/r02f_sweep_n8_deal64.jsonl 2> $O/r02f_sweep_c.err; cat $O/r02f_sweep_n8_deal64.jsonl
ICB_DEAL_BLOCK=256 timeout 300 $TR --master-port 29520 $S --configs "8,8,0,384,16,10;16,8,0,384,16,10" > $O/r02f_sweep_n8_deal256.jsonl 2> $O/r02f_sweep_d.err; cat $O/r02f_sweep_n8_deal256.jsonl
ICB_COMM_NO_PRIORITY=1 timeout 300 $TR --master-port 29521 $S --configs "8,8,0,384,16,10;16,8,0,384,16,10" > $O/r02f_sweep_n8_noprio.jsonl 2> $O/r02f_sweep_e.err; cat $O/r02f_sweep_n8_noprio.jsonl
timeout 400 $TR --master-port 29522 bench.py --gpus 8 --workload C4 --steps 20 --warmup 3 > $O/r02f_bench_c4_n8.json 2> $O/r02f_bench_c4_n8.err; tail -c 500 $O/r02f_bench_c4_n8.json; tail -3 $O/r02f_bench_c4_n8.err
