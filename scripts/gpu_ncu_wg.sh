#!/bin/bash
mkdir -p gpurun_out
timeout 300 python scripts/bench_shapes.py 64 "fc1 wgrad" > gpurun_out/plain2.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:umma_wgrad -s 4 -c 1 -o gpurun_out/prof_wg_fc1 -f python scripts/bench_shapes.py 64 "fc1 wgrad" > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/plain2.log | tail -2
