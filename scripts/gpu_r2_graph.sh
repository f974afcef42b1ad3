#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_graph.py -m gpu -x -q -s > gpurun_out/pytest_graph.log 2>&1; echo "pytest rc=$?"; grep -v "^$" gpurun_out/pytest_graph.log | tail -25 | cut -c1-400
