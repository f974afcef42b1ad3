#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_pdl.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/pytest_pdl.log
for mode in 1 0 1 0; do
S4G_PDL=$mode timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_pdl$mode.json 2> gpurun_out/bench_pdl$mode.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_pdl$mode.json')); print('pdl=$mode', round(d['ms_per_step'],3), round(d['value']), d['step_ms'], round(d['e2e']['value']), d['loss'])"
done
