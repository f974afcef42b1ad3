#!/bin/bash
mkdir -p gpurun_out
timeout 120 python scripts/dev_umma.py wgrad > gpurun_out/dev_wg.log 2>&1; echo "wgrad rc=$?"; tail -7 gpurun_out/dev_wg.log
timeout 120 python scripts/dev_umma.py conv > gpurun_out/dev_conv.log 2>&1; echo "conv rc=$?"; grep wgrad gpurun_out/dev_conv.log
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python scripts/bench_shapes.py 64 wgrad > gpurun_out/shapes8.txt 2>&1; tail -1 gpurun_out/shapes8.txt
S4G_WGRAD_CG=1 timeout 300 python scripts/bench_shapes.py 64 wgrad > gpurun_out/shapes8_cg1.txt 2>&1; tail -1 gpurun_out/shapes8_cg1.txt
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench8.json 2> gpurun_out/bench8.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench8.json')); print(d['ms_per_step'], d['value'], d['loss'])"
