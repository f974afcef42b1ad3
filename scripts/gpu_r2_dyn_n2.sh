#!/bin/bash
mkdir -p gpurun_out
N=${N:-2}
for dyn in 1 0 1 0; do
S4G_DYN_TILES=$dyn timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29553 bench.py --gpus $N --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n${N}_dyn$dyn.json 2> gpurun_out/bench_n${N}_dyn$dyn.err; echo "dyn=$dyn rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/bench_n${N}_dyn$dyn.json').read().strip().splitlines()[-1]); print('N=$N dyn=$dyn', round(d['ms_per_step'],3), 'ms', round(d['value']), 'img/s', d['loss'])"
done
