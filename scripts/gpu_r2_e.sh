#!/bin/bash
mkdir -p gpurun_out
timeout 180 python scripts/dev_attn.py 3 64 > gpurun_out/attn.log 2>&1; echo "attn rc=$?"; grep -E "rel|time|identical|Error|error" gpurun_out/attn.log | head -30
for d in 0 64; do S4G_ATTN_DBG=$d timeout 120 python scripts/dev_attn_time.py 12 64 2>&1 | grep "dbg"; done | tee gpurun_out/attn_dbg12.log
timeout 900 python -m pytest tests/test_gpu_model.py tests/test_gpu_ops.py tests/test_gpu_determinism.py -q -m gpu -x > gpurun_out/pytest_h.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_h.log; tail -3 gpurun_out/pytest_h.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/bench_i.json 2> gpurun_out/bench_i.err
python -c "
import json; d=json.load(open('gpurun_out/bench_i.json')); print('seg ms/step', d['ms_per_step'])"
