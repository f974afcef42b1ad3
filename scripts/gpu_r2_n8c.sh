#!/bin/bash
# 8 GPUs: bucket-wise optimizer step and bucket count A/B for the overlapped gradient exchange (seg), then MAE
mkdir -p gpurun_out
N=${N:-8}
run() {  # tag, task, env...
  tag=$1; task=$2; shift 2
  env "$@" timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29553 bench.py --gpus $N --task $task --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n${N}_$tag.json 2> gpurun_out/bench_n${N}_$tag.err; echo "$tag rc=$?"
  python -c "
import json; d=json.loads(open('gpurun_out/bench_n${N}_$tag.json').read().strip().splitlines()[-1]); print('$tag N=$N', round(d['ms_per_step'],3), 'ms', round(d['value']), 'img/s e2e', round(d['e2e']['value']), 'loss', d['loss'])" || tail -5 gpurun_out/bench_n${N}_$tag.err
}
run seg_fused_b5 seg S4G_X=0
run seg_plain_b5 seg S4G_FUSED_STEP=0
run seg_fused_b9 seg S4G_DDP_BUCKETS=9
run seg_fused_b13 seg S4G_DDP_BUCKETS=13
run mae_fused_b9 mae S4G_DDP_BUCKETS=9
