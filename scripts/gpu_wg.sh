#!/bin/bash
mkdir -p gpurun_out
timeout 120 python scripts/dev_umma.py wgrad > gpurun_out/dev_wg.log 2>&1; echo "wgrad rc=$?"; tail -7 gpurun_out/dev_wg.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python scripts/bench_shapes.py 64 wgrad > gpurun_out/shapes14.txt 2>&1; tail -1 gpurun_out/shapes14.txt
grep -E "^(qkv|proj|fc1|fc2|patch|out_conv|post)" gpurun_out/shapes14.txt
for i in 1 2; do
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench14.json 2> gpurun_out/bench14.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench14.json')); print(d['ms_per_step'], d['value'], d['e2e']['value'], d['loss'])"
done
