#!/bin/bash
# round 2: full GPU validation -- the no-alloc test alone in a fresh process, smoke, every GPU test (as the driver runs them: -x),
# the bench line with its side measurements, reference arm
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_zz_alone.log 2>&1; echo "zz alone rc=$?"; tail -3 gpurun_out/pytest_zz_alone.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 2400 python -m pytest tests -m gpu -x -q -s > gpurun_out/pytest_gpu_full.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_full.log
grep -E "strict_pass|passed|failed|FAILED|Error" gpurun_out/pytest_gpu_full.log | cut -c1-250 | tail -24
timeout 1500 python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench rc=$?"
cat gpurun_out/bench_full.json | cut -c1-4500
timeout 600 python bench.py --impl reference --steps 8 --warmup 2 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
cut -c1-400 gpurun_out/bench_ref.json
