#!/bin/bash
# BASELINE configs[4] sort-first on N GPUs of one box: R2B10 x 90 built once by rank 0, broadcast to the other ranks.
# usage: tools/run_c5_multi.sh N [steps]; a watchdog ends the run (exact PID) if the box runs short of host memory.
N=${1:-2}; K=${2:-8}
O=gpurun_out
free -g | head -2 > $O/c5_n${N}_mem.log; nproc >> $O/c5_n${N}_mem.log
export ICB_BUILD_TIMING=1
if [ "$N" = "1" ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541"; fi
timeout 900 $L \
  bench.py --gpus $N --workload C5 --cpu-seconds 0 --steps $K --warmup 3 > $O/c5_n${N}.json 2> $O/c5_n${N}.err &
PID=$!
MINAV=999999999
while kill -0 $PID 2>/dev/null; do
  AV=$(awk '/MemAvailable/{print $2}' /proc/meminfo)
  [ "$AV" -lt "$MINAV" ] && MINAV=$AV
  if [ "$AV" -lt 6000000 ]; then echo "watchdog: MemAvailable $AV kB, stopping" >> $O/c5_n${N}_mem.log; kill $PID; break; fi
  sleep 1
done
wait $PID; echo "rc=$? min MemAvailable kB=$MINAV" >> $O/c5_n${N}_mem.log
cat $O/c5_n${N}_mem.log; tail -c 1200 $O/c5_n${N}.json; grep -v Warning $O/c5_n${N}.err | tail -25
