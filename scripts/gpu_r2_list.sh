#!/bin/bash
# ncu launch list (time only) of one training step on the final round-2 code
mkdir -p gpurun_out
timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2600 --csv --log-file gpurun_out/launches_r02.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/ncu_list.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/launches_r02.csv; tail -1 gpurun_out/plain.log | cut -c1-200
timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-extras 2>/dev/null | cut -c1-400
