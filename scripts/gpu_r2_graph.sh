#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_graph.py -m gpu -x -q -s > gpurun_out/pytest_graph.log 2>&1; echo "pytest rc=$?"; grep -v "^$" gpurun_out/pytest_graph.log | tail -12 | cut -c1-300
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_graph.json 2> gpurun_out/bench_graph.err; echo "bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/bench_graph.json').read().strip().splitlines()[-1]); print(round(d['ms_per_step'],3), round(d['value'])); print(d['extra']['cuda_graph']); print({k:round(v['frac'],3) for k,v in d['roofline']['family'].items() if isinstance(v,dict)}, d['roofline']['family'].get('error'))"
