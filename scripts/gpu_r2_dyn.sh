#!/bin/bash
# dynamic tile scheduler: operator / model parity, then A/B against the static partition at N=1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_ops.py tests/test_gpu_model.py tests/test_gpu_determinism.py tests/test_gpu_mae.py -m gpu -q -x > gpurun_out/pytest_dyn.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_dyn.log | cut -c1-300
for dyn in 1 0 1 0; do
S4G_DYN_TILES=$dyn timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/bench_dyn$dyn.json 2> gpurun_out/bench_dyn$dyn.err; echo "dyn=$dyn rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/bench_dyn$dyn.json').read().strip().splitlines()[-1]); print('dyn=$dyn', round(d['ms_per_step'],3), 'ms', round(d['value']), 'img/s', d['loss'])"
done
