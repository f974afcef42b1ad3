#!/bin/bash
# round 2, call A: GPU tests after the determinism / ADVICE changes, conditioning diagnostics, loss curves, a bench line
set -x
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 1500 python -m pytest tests -m gpu -q -s --deselect tests/test_gpu_train.py > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
grep -E "strict_pass|passed|failed|FAILED|rc=" gpurun_out/pytest_gpu.log | tail -30
for d in seg depth; do timeout 300 python tests/tools/dev_wellcond.py $d 4 50 > gpurun_out/wellcond_$d.log 2>&1; head -4 gpurun_out/wellcond_$d.log; done
timeout 900 python tests/tools/dev_curve.py seg depth cls > gpurun_out/curves.log 2>&1; grep -E "max dev" gpurun_out/curves.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_a.json 2> gpurun_out/bench_a.err; echo "bench rc=$?"
cat gpurun_out/bench_a.json
