#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 120 python scripts/dev_umma.py lin > gpurun_out/dev_lin.log 2>&1; echo "lin rc=$?"
cat gpurun_out/dev_lin.log | tail -12
timeout 120 python scripts/dev_umma.py conv > gpurun_out/dev_conv.log 2>&1; echo "conv rc=$?"
cat gpurun_out/dev_conv.log | tail -16
nvidia-smi --query-gpu=name,memory.used --format=csv
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python scripts/bench_shapes.py 64 > gpurun_out/shapes4.txt 2>&1; tail -1 gpurun_out/shapes4.txt
S4G_FPROP_CG=1 timeout 300 python scripts/bench_shapes.py 64 > gpurun_out/shapes4_cg1.txt 2>&1; tail -1 gpurun_out/shapes4_cg1.txt
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench4.json 2> gpurun_out/bench4.err; echo "bench rc=$?"
cat gpurun_out/bench4.json
