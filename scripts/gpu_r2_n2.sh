#!/bin/bash
# 2 GPUs: SyncBN peer exchange checks, train_epoch with rank-0-only evaluation, seg + depth bench at N=2
mkdir -p gpurun_out
{
echo "# round 2, 2 x B200 (one box): multi-GPU checks (scripts/check_syncbn_n2.py, scripts/check_train_n2.py via tests/test_gpu_multi.py)"
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q -s 2>&1 | tail -15
for s in check_syncbn_n2 check_train_n2; do
  echo "## torchrun --nproc-per-node 2 scripts/$s.py"
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 scripts/$s.py 2>&1 | grep -v "^W\|^\*\*\*\|OMP_NUM" | tail -12
done
} > gpurun_out/r02_multi_checks.txt 2>&1
cat gpurun_out/r02_multi_checks.txt
[ -n "$SKIP_BENCH" ] && exit 0
for task in seg depth; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29552 bench.py --gpus 2 --task $task --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n2_$task.json 2> gpurun_out/bench_n2_$task.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_n2_$task.json').read().strip().splitlines()[-1]); print('$task N=2', round(d['ms_per_step'],3), 'ms', round(d['value']), 'img/s e2e', round(d['e2e']['value']))"
done
