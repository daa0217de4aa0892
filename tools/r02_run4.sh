#!/bin/bash
# round 2, fourth GPU call (eight B200): group tests at 4 ranks, sort-first sweeps and bench lines at N=8 and N=4
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 python -m pytest tests/test_group.py tests/test_woodcock_grid.py tests/test_woodcock_sphere.py -m gpu -x -q > $O/r02d_pytest.log 2>&1; tail -3 $O/r02d_pytest.log
timeout 400 $TR --nproc-per-node 8 --master-port 29517 tools/sweep_scale.py \
  --configs "4,2,0,384,16,10;4,2,1,384,16,10;8,4,0,384,16,10;8,2,0,384,16,10;8,4,1,384,16,10;4,4,0,384,16,10;4,1,0,384,16,10;2,1,0,384,16,10;2,2,1,384,16,10;8,8,0,384,16,10;1,1,0,384,16,10;1,1,1,384,16,10" > $O/r02d_sweep_n8.jsonl 2> $O/r02d_sweep_n8.err
cat $O/r02d_sweep_n8.jsonl; tail -5 $O/r02d_sweep_n8.err
timeout 300 $TR --nproc-per-node 8 --master-port 29518 bench.py --gpus 8 --steps 20 --warmup 3 > $O/r02d_bench_n8.json 2> $O/r02d_bench_n8.err
tail -c 600 $O/r02d_bench_n8.json; tail -5 $O/r02d_bench_n8.err
timeout 300 $TR --nproc-per-node 4 --master-port 29519 tools/sweep_scale.py \
  --configs "4,2,0,384,16,10;4,2,1,384,16,10;8,4,0,384,16,10;4,4,0,384,16,10;4,1,0,384,16,10;2,1,0,384,16,10" > $O/r02d_sweep_n4.jsonl 2> $O/r02d_sweep_n4.err
cat $O/r02d_sweep_n4.jsonl; tail -5 $O/r02d_sweep_n4.err
timeout 300 $TR --nproc-per-node 4 --master-port 29520 bench.py --gpus 4 --steps 20 --warmup 3 > $O/r02d_bench_n4.json 2> $O/r02d_bench_n4.err
tail -c 600 $O/r02d_bench_n4.json; tail -5 $O/r02d_bench_n4.err
