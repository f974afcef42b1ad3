#!/bin/bash
# One gpurun call: GPU tests, bench line, per-shape timings, ncu launch list, one full ncu capture of the top kernel.
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json
timeout 600 python bench.py --impl reference --steps 8 --warmup 2 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
cat gpurun_out/bench_ref.json
timeout 600 python scripts/bench_shapes.py 64 > gpurun_out/shapes.txt 2>&1; echo "shapes rc=$?"
tail -1 gpurun_out/shapes.txt
timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
timeout 300 python scripts/bench_shapes.py 64 "head112 fwd" > gpurun_out/plain2.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:umma_fprop -s 4 -c 1 -o gpurun_out/prof_fprop_head112 -f python scripts/bench_shapes.py 64 "head112 fwd" > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
