#!/bin/bash
# attention v4 integrated: tests, loss curves (depth bound), conditioning at B=64, bench
set -x
mkdir -p gpurun_out
timeout 120 python scripts/dev_attn.py 64 > gpurun_out/attn.log 2>&1; grep -E "time" gpurun_out/attn.log
timeout 1500 python -m pytest tests -m gpu -q -s --deselect tests/test_gpu_train.py --deselect tests/test_gpu_wellcond.py > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
grep -E "strict_pass|passed|failed|FAILED|rc=" gpurun_out/pytest_gpu.log | tail -12
for d in seg depth; do timeout 400 python tests/tools/dev_wellcond.py $d 64 30 > gpurun_out/wellcond64_$d.log 2>&1; grep -v "^   l2 " gpurun_out/wellcond64_$d.log; done
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_f.json 2> gpurun_out/bench_f.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_f.json')); print('seg ms/step', d['ms_per_step'], 'img/s', d['value'], 'e2e', d['e2e']['value'])"
S4G_ATTN=mma timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_f_mma.json 2> gpurun_out/bench_f_mma.err
python -c "
import json; d=json.load(open('gpurun_out/bench_f_mma.json')); print('seg (old mma.sync attention) ms/step', d['ms_per_step'])"
