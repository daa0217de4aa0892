#!/bin/bash
# R2B10 x 90 (BASELINE configs[4], 251.7 M records): tests/test_gpu_parity.py::test_c5_parity_at_headline_size -- >= 600
# positions located by icb_locate (plain, walk, exact) against the reference's brute-force loop over all records.
# The host holds the 71.5 GB of records plus the builder's arrays; a watchdog ends the run (exact PID) if memory runs short.
O=gpurun_out
free -g | head -2 > $O/r03_c5_parity_mem.log
ICB_TEST_C5=1 timeout ${1:-2400} python -m pytest tests/test_gpu_parity.py -m gpu -k test_c5_parity_at_headline_size -v -rs --durations=3 > $O/r03_c5_parity.log 2>&1 &
PID=$!
MINAV=999999999
while kill -0 $PID 2>/dev/null; do
  AV=$(awk '/MemAvailable/{print $2}' /proc/meminfo)
  [ "$AV" -lt "$MINAV" ] && MINAV=$AV
  if [ "$AV" -lt 6000000 ]; then echo "watchdog: MemAvailable $AV kB, stopping" >> $O/r03_c5_parity_mem.log; kill $PID; break; fi
  sleep 1
done
wait $PID; echo "rc=$? min MemAvailable kB=$MINAV nproc=$(nproc)" >> $O/r03_c5_parity_mem.log
cat $O/r03_c5_parity_mem.log; tail -15 $O/r03_c5_parity.log
