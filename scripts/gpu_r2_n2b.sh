#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 scripts/check_mae_n2.py 2>&1 | grep -v "^W\|^\*\*\*\|OMP_NUM" | tail -25 | tee gpurun_out/check_mae_n2.txt
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29553 bench.py --gpus 2 --task mae --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n2_mae.json 2> gpurun_out/bench_n2_mae.err; echo "mae rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/bench_n2_mae.json').read().strip().splitlines()[-1]); print('mae N=2', round(d['ms_per_step'],3), 'ms', round(d['value']), 'img/s e2e', round(d['e2e']['value']), 'loss', d['loss'])" || tail -5 gpurun_out/bench_n2_mae.err
