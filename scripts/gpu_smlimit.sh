#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
run() { tag=$1; shift
env "$@" timeout 600 $TR --master-port 29522 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_n2_$tag.json 2> gpurun_out/bench_n2_$tag.err; echo "bench n2 $tag rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_n2_$tag.json')); print('$tag', d['n_gpus'], round(d['ms_per_step'],3), round(d['value']), round(d['e2e']['value']), d['loss'])"
}
run base A=1
run sm140_c8 S4G_SM_LIMIT=140 NCCL_MAX_CTAS=8
run sm132_c16 S4G_SM_LIMIT=132 NCCL_MAX_CTAS=16
run sm132_c8 S4G_SM_LIMIT=132 NCCL_MAX_CTAS=8
run sm124_c24 S4G_SM_LIMIT=124 NCCL_MAX_CTAS=24
run sm140_def S4G_SM_LIMIT=140
