#!/bin/bash
mkdir -p gpurun_out
for bn in 256 192 128; do
  S4G_FPROP_BN=$bn timeout 300 python scripts/bench_shapes.py 64 > gpurun_out/shapes5_bn$bn.txt 2>&1; tail -1 gpurun_out/shapes5_bn$bn.txt
done
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench5.json 2> gpurun_out/bench5.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench5.json')); print(d['ms_per_step'], d['value'])"
