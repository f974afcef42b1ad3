#!/bin/bash
mkdir -p gpurun_out
for cfg in "0 0" "1 0" "0 1" "1 1"; do
  set -- $cfg
  S4G_FUSE_GELU=$1 S4G_FUSE_DGELU=$2 timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_f$1$2.json 2> gpurun_out/bench_f.err
  python -c "
import json,sys; d=json.load(open('gpurun_out/bench_f$1$2.json')); print('gelu=$1 dgelu=$2', d['ms_per_step'], d['loss'])"
done
