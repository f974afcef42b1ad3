#!/bin/bash
mkdir -p gpurun_out
for i in 1 2 3 4; do
S4G_BENCH_TRACE=1 timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_tr$i.json 2> gpurun_out/bench_tr$i.err; echo "bench rc=$?"
grep "^trace" gpurun_out/bench_tr$i.err
python -c "
import json; d=json.load(open('gpurun_out/bench_tr$i.json')); print(round(d['ms_per_step'],2), round(d['value']), d['step_ms'], round(d['e2e']['value']), d['e2e']['step_ms'])"
done
