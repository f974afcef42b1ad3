#!/bin/bash
# 2-GPU call: peer-memory SyncBN exchange check, then bench at N=2 with the peer kernel and with the NCCL callback.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29521 scripts/check_syncbn_n2.py > gpurun_out/check_syncbn_n2.log 2>&1; echo "check rc=$?"
grep -v "^W\|^\*\*\*" gpurun_out/check_syncbn_n2.log | tail -12
for mode in peer nccl; do
S4G_SYNCBN=$mode timeout 600 $TR --master-port 29522 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_n2_$mode.json 2> gpurun_out/bench_n2_$mode.err; echo "bench n2 $mode rc=$?"
tail -2 gpurun_out/bench_n2_$mode.err
python -c "
import json; d=json.load(open('gpurun_out/bench_n2_$mode.json')); print('$mode', d['n_gpus'], d['ms_per_step'], d['value'], d['e2e'], d['loss'])"
done
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench15.json 2> gpurun_out/bench15.err; echo "bench n1 rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench15.json')); print('n1', d['ms_per_step'], d['value'], d['e2e'], d['roofline']['frac'])"
