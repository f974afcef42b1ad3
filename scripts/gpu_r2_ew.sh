#!/bin/bash
# HBM-pass tuning (pack, split-K reduce, bn_apply, bilinear backward): parity, then the bench twice
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ops.py tests/test_gpu_model.py tests/test_gpu_determinism.py tests/test_gpu_fullsize.py tests/test_gpu_mae.py -m gpu -q -x > gpurun_out/pytest_ew.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_ew.log | cut -c1-300
for v in 1 2; do
timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/bench_ew_$v.json 2> gpurun_out/bench_ew_$v.err; echo "$v rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/bench_ew_$v.json').read().strip().splitlines()[-1]); print('run $v', round(d['ms_per_step'],3), 'ms', round(d['value']), 'img/s', d['loss'], d['gpu_launches'])"
done

timeout 300 python scripts/bench_shapes.py 64 2>&1 | grep -E "out_conv(14|28|56) wgrad|proj wgrad|TOTAL"
