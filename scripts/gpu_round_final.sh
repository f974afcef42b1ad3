#!/bin/bash
# Final round-1 call: smoke, GPU tests, bench line, reference arm, per-shape timings, ncu launch list, one ncu --set full
# capture of the fill-bound token GEMM (fc1 forward, CTA pairs).
set -x
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json
timeout 600 python bench.py --impl reference --steps 8 --warmup 2 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
cat gpurun_out/bench_ref.json
for task in depth cls; do
timeout 600 python bench.py --task $task --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_$task.json 2> gpurun_out/bench_$task.err; echo "$task rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_$task.json')); print('$task', round(d['ms_per_step'],3), round(d['value']), round(d['e2e']['value']))"
done
timeout 600 python scripts/bench_shapes.py 64 > gpurun_out/shapes.txt 2>&1; echo "shapes rc=$?"
tail -1 gpurun_out/shapes.txt
timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
timeout 300 python scripts/bench_shapes.py 64 "fc1 fwd" > gpurun_out/plain2.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:umma_fprop -s 4 -c 1 -o gpurun_out/prof_fprop_fc1 -f python scripts/bench_shapes.py 64 "fc1 fwd" > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
