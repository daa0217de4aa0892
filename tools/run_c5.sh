#!/bin/bash
# BASELINE configs[4] on one B200: R2B10 x 90 replicated grid.  The host side holds the 71.5 GB of records plus the
# builder's arrays (~160 GB at the peak); a watchdog ends the run (exact PID) if the box runs short of memory.
# The CPU-baseline leg (C5_CPU_SECONDS, default 60) locates ~50 positions with the reference loop and compares them with icb_locate (parity_checked).
O=gpurun_out
free -g | head -2 > $O/c5_mem.log
timeout ${C5_TIMEOUT:-600} python bench.py --workload C5 --cpu-seconds ${C5_CPU_SECONDS:-60} --steps 8 --warmup 3 > $O/c5.json 2> $O/c5.err &
PID=$!
MINAV=999999999
while kill -0 $PID 2>/dev/null; do
  AV=$(awk '/MemAvailable/{print $2}' /proc/meminfo)
  [ "$AV" -lt "$MINAV" ] && MINAV=$AV
  if [ "$AV" -lt 6000000 ]; then echo "watchdog: MemAvailable $AV kB, stopping" >> $O/c5_mem.log; kill $PID; break; fi
  sleep 1
done
wait $PID; echo "rc=$? min MemAvailable kB=$MINAV" >> $O/c5_mem.log
cat $O/c5_mem.log; tail -c 1500 $O/c5.json; tail -3 $O/c5.err
