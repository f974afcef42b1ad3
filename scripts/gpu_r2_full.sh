#!/bin/bash
# round 2: full GPU validation -- smoke, every GPU test, the bench line with its side measurements, reference arm
set -x
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 2400 python -m pytest tests -m gpu -q -s > gpurun_out/pytest_gpu_full.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_full.log
grep -E "strict_pass|loss-curve|passed|failed|FAILED|rc=" gpurun_out/pytest_gpu_full.log | cut -c1-250 | tail -30
timeout 1500 python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench rc=$?"
cat gpurun_out/bench_full.json | cut -c1-3000
timeout 600 python bench.py --impl reference --steps 8 --warmup 2 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
