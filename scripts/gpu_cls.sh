#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --task cls --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_cls.json 2> gpurun_out/bench_cls.err; echo "cls rc=$?"; tail -3 gpurun_out/bench_cls.err
python -c "
import json; d=json.load(open('gpurun_out/bench_cls.json')); print('cls', d['ms_per_step'], d['value'], d['model_tflops_per_gpu'], d['loss'], d['e2e'])"
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench9.json 2> gpurun_out/bench9.err; echo "seg rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench9.json')); print('seg', d['ms_per_step'], d['value'], d['loss'])"
