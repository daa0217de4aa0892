#!/bin/bash
# four B200s: the group tests (1/2/4 ranks), BASELINE configs[4] and configs[3], and the time series with the in-place swap
O=gpurun_out
timeout 600 python -m pytest tests/test_group.py -m gpu -x -q > $O/r02l_pytest_group.log 2>&1; tail -3 $O/r02l_pytest_group.log
bash tools/run_configs.sh 4
L="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29543"
ICB_NO_LINES_COPY=1 timeout 600 $L bench.py --gpus 4 --workload C4 --cpu-seconds 0 --steps 24 --warmup 3 --timesteps 24 > $O/c4_n4_inplace.json 2> $O/c4_n4_inplace.err
tail -c 600 $O/c4_n4_inplace.json
