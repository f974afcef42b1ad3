#!/bin/bash
mkdir -p gpurun_out
timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
timeout 1200 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active --clock-control none -c 1800 --csv --log-file gpurun_out/step_metrics.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_metrics.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/step_metrics.csv
