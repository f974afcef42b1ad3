#!/bin/bash
mkdir -p gpurun_out
timeout 300 python scripts/bench_shapes.py 64 > gpurun_out/shapes7.txt 2>&1; tail -1 gpurun_out/shapes7.txt
grep -E "^(qkv|proj|fc1|fc2) (fwd|dgrad)" gpurun_out/shapes7.txt
timeout 120 python scripts/prof_attn.py > gpurun_out/attn_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_ -s 2 -c 2 -o gpurun_out/prof_attn -f python scripts/prof_attn.py > gpurun_out/ncu_attn.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/attn_plain.log | tail -5; ls -la gpurun_out/*.ncu-rep
