#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_augment.py -m gpu -q -s -x > gpurun_out/pytest_aug.log 2>&1; echo "pytest rc=$?"; grep -v "^$" gpurun_out/pytest_aug.log | tail -30 | cut -c1-400
timeout 300 python scripts/bench_augment.py 64 2>&1 | tail -3 | tee gpurun_out/bench_augment.txt
