#!/bin/bash
# round-2 profiles: per-shape GEMM table, attention ncu --set full (fwd, delta, bwd), launch list of one step
mkdir -p gpurun_out
timeout 300 python scripts/bench_shapes.py 64 > gpurun_out/r02_shapes.txt 2>&1; tail -3 gpurun_out/r02_shapes.txt
timeout 120 python scripts/prof_attn.py > gpurun_out/attn_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_ -s 3 -c 3 -o gpurun_out/prof_attn_r02 -f python scripts/prof_attn.py > gpurun_out/ncu_attn.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/prof_attn_r02.ncu-rep
