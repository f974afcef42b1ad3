#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_wellcond.py tests/test_gpu_train.py -m gpu -q -s > gpurun_out/pytest_i.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_i.log
timeout 900 python bench.py > gpurun_out/bench_i.json 2> gpurun_out/bench_i.err; echo "bench rc=$?"; cat gpurun_out/bench_i.json
timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/plain.log 2>&1 &&
timeout 1200 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active --clock-control none -c 2600 --csv --log-file gpurun_out/step_metrics.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/ncu_metrics.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/step_metrics.csv
