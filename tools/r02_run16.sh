#!/bin/bash
# one B200: all GPU tests on the build with the direct node load (v2 made the default library)
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $O/r02o_pytest.log 2>&1; tail -3 $O/r02o_pytest.log
