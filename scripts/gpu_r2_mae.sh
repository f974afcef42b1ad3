#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_mae.py -m gpu -q -s -x > gpurun_out/pytest_mae.log 2>&1; echo "pytest rc=$?"; grep -v "^$" gpurun_out/pytest_mae.log | tail -40 | cut -c1-600
for b in 64 256; do
timeout 600 python bench.py --task mae --batch $b --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_mae_$b.json 2> gpurun_out/bench_mae_$b.err; echo "mae b=$b rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/bench_mae_$b.json').read().strip().splitlines()[-1]); print('mae B=$b', round(d['ms_per_step'],3), 'ms', round(d['value']), 'img/s e2e', round(d['e2e']['value']), 'TF', round(d['model_tflops_per_gpu'],1), 'launches', d['gpu_launches'], 'loss', d['loss'])" || tail -5 gpurun_out/bench_mae_$b.err
done
