#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29521 scripts/check_syncbn_n2.py > gpurun_out/check_syncbn_n2.log 2>&1; echo "check rc=$?"
grep -v "^W\|^\*\*\*" gpurun_out/check_syncbn_n2.log | grep -v "^\[rank\|^  \|^Traceback\|^---\|^===" | tail -8
run() { tag=$1; shift
env "$@" timeout 600 $TR --master-port 29522 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_n2_$tag.json 2> gpurun_out/bench_n2_$tag.err; echo "bench n2 $tag rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_n2_$tag.json')); print('$tag', d['n_gpus'], round(d['ms_per_step'],3), round(d['value']), round(d['e2e']['value']), d['loss'])"
}
run base A=1
run nooverlap S4G_DDP_OVERLAP=0
run cta4 NCCL_MAX_CTAS=4
run cta32 NCCL_MIN_CTAS=32
