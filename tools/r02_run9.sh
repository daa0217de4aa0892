#!/bin/bash
# re-entry check of HEAD on one B200: GPU tests, smoke, the default bench line and the reference arm
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $O/r02i_pytest.log 2>&1; tail -3 $O/r02i_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/r02i_smoke.log 2>&1; tail -2 $O/r02i_smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 > $O/r02i_bench_n1.json 2> $O/r02i_bench_n1.err; tail -c 600 $O/r02i_bench_n1.json; tail -2 $O/r02i_bench_n1.err
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > $O/r02i_bench_ref.json 2> $O/r02i_bench_ref.err; tail -c 400 $O/r02i_bench_ref.json
