#!/bin/bash
# round 2, one B200: all GPU tests after the plain-compositing / computed-logf changes, bench lines of every kernel, logf table A/B
O=gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > $O/r02g_pytest.log 2>&1; tail -4 $O/r02g_pytest.log
B="python bench.py --cpu-seconds 0"
timeout 600 python bench.py --steps 20 --warmup 3 > $O/r02g_bench_n1.json 2> $O/r02g_bench_n1.err; tail -c 300 $O/r02g_bench_n1.json
for v in "--sampler prism:prism" "--integrator woodcock_grid:wgrid" "--integrator woodcock_sphere:wsphere" "--integrator woodcock --steps 4:wae"; do
  a="${v%%:*}"; n="${v##*:}"
  timeout 600 $B $a > $O/r02g_bench_$n.json 2> $O/r02g_bench_$n.err; tail -c 200 $O/r02g_bench_$n.json; tail -2 $O/r02g_bench_$n.err
  ICB_LOG_TABLE=1 timeout 600 $B $a > $O/r02g_bench_${n}_table.json 2> /dev/null
done
ICB_BUILD_TIMING=1 timeout 300 $B --steps 2 --min-region 0 2> $O/r02g_build_timing.log > /dev/null; grep buildGrid $O/r02g_build_timing.log
