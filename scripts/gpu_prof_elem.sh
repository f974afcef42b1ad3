#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench2.json 2> gpurun_out/bench2.err; echo "bench rc=$?"
cat gpurun_out/bench2.json
timeout 600 python scripts/bench_shapes.py 64 > gpurun_out/shapes2.txt 2>&1; tail -1 gpurun_out/shapes2.txt
S4G_FPROP_BN=256 timeout 600 python scripts/bench_shapes.py 64 > gpurun_out/shapes2_bn256.txt 2>&1; tail -1 gpurun_out/shapes2_bn256.txt
S4G_FPROP_BN=128 timeout 600 python scripts/bench_shapes.py 64 > gpurun_out/shapes2_bn128.txt 2>&1; tail -1 gpurun_out/shapes2_bn128.txt
timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none -k regex:"bn_bwd_apply|bn_bwd_reduce|seghead|bilinear2x_bwd|bilinear2x_fwd|layernorm_bwd|dgelu|gelu_kernel|wgrad_reduce_linear" -s 680 -c 24 -o gpurun_out/prof_elem -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_elem.log 2>&1
echo "ncu rc=$?"
ls -la gpurun_out
