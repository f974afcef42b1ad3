#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_model.py -k prefetch -x -q 2>&1 | tail -2
for mode in smi nosmi; do
for i in 1 2 3 4 5 6; do
if [ $mode = nosmi ]; then export S4G_BENCH_NOSMI=1; fi
S4G_BENCH_TRACE=1 timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_tr$i.json 2> gpurun_out/bench_tr$i.err
grep "^trace from_host=True" gpurun_out/bench_tr$i.err | cut -c1-120
python -c "
import json; d=json.load(open('gpurun_out/bench_tr$i.json')); print('$mode', round(d['ms_per_step'],2), round(d['value']), round(d['e2e']['value']), d['e2e']['step_ms'])"
done
done
