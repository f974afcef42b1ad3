#!/bin/bash
# 8 GPUs: depth (config 3), MAE pretraining (config 5) and seg (config 2) bench lines
mkdir -p gpurun_out
N=${N:-8}
for task in depth mae seg; do
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29553 bench.py --gpus $N --task $task --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n${N}_$task.json 2> gpurun_out/bench_n${N}_$task.err; echo "$task rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/bench_n${N}_$task.json').read().strip().splitlines()[-1]); print('$task N=$N', round(d['ms_per_step'],3), 'ms', round(d['value']), 'img/s e2e', round(d['e2e']['value']), 'loss', d['loss'], d['clocks'])" || tail -5 gpurun_out/bench_n${N}_$task.err
done
