#!/bin/bash
# round 2, first GPU call (one B200): GPU tests, default bench line, tail diagnosis
O=gpurun_out
python -m pytest tests -m gpu -x -q > $O/r02a_pytest.log 2>&1; tail -5 $O/r02a_pytest.log
python bench.py > $O/r02a_bench_n1.json 2> $O/r02a_bench_n1.err; tail -c 600 $O/r02a_bench_n1.json; tail -3 $O/r02a_bench_n1.err
python tools/diag_tail.py > $O/r02a_diag_tail.log 2>&1; tail -12 $O/r02a_diag_tail.log
