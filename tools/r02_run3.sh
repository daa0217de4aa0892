#!/bin/bash
# round 2, third GPU call (two B200): all GPU tests (2-rank group cases included), sort-first sweep and bench at N=2
O=gpurun_out
python -m pytest tests -m gpu -x -q > $O/r02c_pytest.log 2>&1; tail -5 $O/r02c_pytest.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 tools/sweep_scale.py \
  --configs "4,2,0,384,16,10;4,2,1,384,16,10;4,4,1,384,16,10;8,4,1,384,16,10;8,4,0,384,16,10;1,1,0,384,16,10;1,1,1,384,16,10" > $O/r02c_sweep_n2.jsonl 2> $O/r02c_sweep_n2.err
cat $O/r02c_sweep_n2.jsonl; tail -5 $O/r02c_sweep_n2.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 2 --steps 20 --warmup 3 > $O/r02c_bench_n2.json 2> $O/r02c_bench_n2.err
tail -c 1200 $O/r02c_bench_n2.json; tail -5 $O/r02c_bench_n2.err
