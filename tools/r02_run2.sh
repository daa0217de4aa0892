#!/bin/bash
# round 2, second GPU call (one B200): split rays + group-of-one tests, split-parameter sweep
O=gpurun_out
python -m pytest tests -m gpu -x -q > $O/r02b_pytest.log 2>&1; tail -5 $O/r02b_pytest.log
for cfg in "0 15" "384 15" "384 14" "384 16" "1024 14" "1024 13"; do
  set -- $cfg
  ICB_SPLIT_HEAVY=$1 ICB_SPLIT_COST_LOG2=$2 python bench.py --steps 10 --cpu-seconds 0 > $O/r02b_split_$1_$2.json 2> $O/r02b_split_$1_$2.err
  python - <<PY
import json
try:
    d=json.load(open("$O/r02b_split_$1_$2.json"))
    print("split heavy=$1 costlog2=$2: ms/step %.3f latency %.3f e2e %.3f kernel_alone %.3f samples %.4g" % (d["ms_per_step"], d["latency_ms"], d["e2e"]["ms_per_step"], d["roofline"]["kernel_ms_alone"], d["config"]["samples_per_step"]))
except Exception as e:
    print("split $1 $2 failed", e)
PY
done
python tools/diag_tail.py > $O/r02b_diag_tail.log 2>&1; tail -12 $O/r02b_diag_tail.log
