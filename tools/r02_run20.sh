#!/bin/bash
# round-end check on one B200: GPU tests, smoke(), the default bench line, the reference arm, the other kernels' lines
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $O/r02s_pytest.log 2>&1; tail -2 $O/r02s_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/r02s_smoke.log 2>&1; tail -1 $O/r02s_smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 > $O/r02s_bench_n1.json 2> $O/r02s_bench_n1.err; tail -c 400 $O/r02s_bench_n1.json
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > $O/r02s_bench_ref.json 2> $O/r02s_bench_ref.err; tail -c 300 $O/r02s_bench_ref.json
timeout 300 python bench.py --sampler prism --cpu-seconds 0 > $O/r02s_bench_prism.json 2>/dev/null
timeout 300 python bench.py --integrator woodcock_grid --cpu-seconds 0 > $O/r02s_bench_wgrid.json 2>/dev/null
timeout 300 python bench.py --integrator woodcock_sphere --cpu-seconds 0 > $O/r02s_bench_wsphere.json 2>/dev/null
timeout 300 python bench.py --integrator woodcock --cpu-seconds 0 --steps 5 > $O/r02s_bench_wae.json 2>/dev/null
timeout 300 python bench.py --in-flight 1 --cpu-seconds 0 > $O/r02s_bench_inflight1.json 2>/dev/null
for f in prism wgrid wsphere wae inflight1; do python - <<P
import json
try:
    b=json.loads(open("$O/r02s_bench_$f.json").read().strip().splitlines()[-1]); print("$f", round(b["ms_per_step"],2), round(b.get("latency_ms") or 0,2))
except Exception as e: print("$f failed", e)
P
done
