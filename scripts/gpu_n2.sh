#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench n2 rc=$?"
tail -3 gpurun_out/bench_n2.err
python -c "
import json; d=json.load(open('gpurun_out/bench_n2.json')); print(d['n_gpus'], d['ms_per_step'], d['value'], d['e2e']['value'], d['loss'])"
timeout 600 python bench.py --task depth --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_depth.json 2> gpurun_out/bench_depth.err; echo "depth rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_depth.json')); print('depth', d['ms_per_step'], d['value'], d['loss'])"
