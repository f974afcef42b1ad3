#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_harness.py tests/test_gpu_model.py -k "harness or prefetch or resume or interchanges" -x -q > gpurun_out/pytest_harness.log 2>&1; echo "harness rc=$?"
tail -15 gpurun_out/pytest_harness.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench16.json 2> gpurun_out/bench16.err; echo "bench rc=$?"
cat gpurun_out/bench16.json
