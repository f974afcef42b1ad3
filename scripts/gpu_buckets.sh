#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
for nb in 5 13 2 9; do
S4G_DDP_BUCKETS=$nb timeout 600 $TR --master-port 29522 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_n2_b$nb.json 2> gpurun_out/bench_n2_b$nb.err; echo "rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_n2_b$nb.json')); print('buckets $nb', round(d['ms_per_step'],3), round(d['value']), round(d['e2e']['value']), d['loss'])"
done
