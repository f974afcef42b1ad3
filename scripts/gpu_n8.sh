#!/bin/bash
# 8-GPU call: SyncBN peer exchange check at world 8, then the bench line at N=8 and N=4.
mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 $TR --nproc-per-node $N --master-port 29521 scripts/check_syncbn_n2.py > gpurun_out/check_syncbn_n$N.log 2>&1; echo "check rc=$?"
grep "vs\|ok\|OK\|Assert" gpurun_out/check_syncbn_n$N.log | tail -8
for n in $N 4; do
timeout 600 $TR --nproc-per-node $n --master-port 29522 bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; echo "bench n$n rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_n$n.json')); print('n$n', d['n_gpus'], round(d['ms_per_step'],3), round(d['value']), d['step_ms'], d['e2e'], d['loss'])"
done
