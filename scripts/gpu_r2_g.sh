#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_wellcond.py -q -s -m gpu > gpurun_out/pytest_wellcond.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_wellcond.log
grep -E "trained state|passed|failed|rc=|Error" gpurun_out/pytest_wellcond.log
timeout 1200 python tests/tools/dev_curve.py seg depth cls > gpurun_out/curves.log 2>&1; cat gpurun_out/curves.log | cut -c1-400
