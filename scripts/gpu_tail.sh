#!/bin/bash
mkdir -p gpurun_out
timeout 120 python scripts/dev_umma.py lin > gpurun_out/dev_lin.log 2>&1; echo "lin rc=$?"; tail -8 gpurun_out/dev_lin.log
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python scripts/bench_shapes.py 64 > gpurun_out/shapes6.txt 2>&1; tail -1 gpurun_out/shapes6.txt
grep -E "^(qkv|proj|fc1|fc2) (fwd|dgrad)" gpurun_out/shapes6.txt
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench6.json 2> gpurun_out/bench6.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench6.json')); print(d['ms_per_step'], d['value'], d['loss'])"
