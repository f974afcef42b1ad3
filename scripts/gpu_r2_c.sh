#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 120 python scripts/prof_attn.py > gpurun_out/plain_attn.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_bwd_umma -s 2 -c 1 -o gpurun_out/prof_attn_bwd4 -f python scripts/prof_attn.py > gpurun_out/ncu_attn.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_attn.log
