#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_model.py tests/test_gpu_fullsize.py tests/test_gpu_determinism.py -q -m gpu -x > gpurun_out/pytest_h.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_h.log; tail -3 gpurun_out/pytest_h.log
for v in fused pass; do
S4G_GELU_BWD=$v timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_h_$v.json 2> gpurun_out/bench_h_$v.err
python -c "
import json; d=json.load(open('gpurun_out/bench_h_$v.json')); print('$v seg ms/step', d['ms_per_step'])"
done
