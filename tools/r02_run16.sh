#!/bin/bash
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $O/r02o_pytest.log 2>&1; tail -3 $O/r02o_pytest.log
