#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 120 python scripts/dev_attn.py 1 > gpurun_out/attn_b1.log 2>&1; echo "attn B=1 rc=$?"; tail -12 gpurun_out/attn_b1.log
timeout 180 python scripts/dev_attn.py 3 64 > gpurun_out/attn.log 2>&1; echo "attn rc=$?"; tail -30 gpurun_out/attn.log
timeout 120 python scripts/prof_attn.py > gpurun_out/plain_attn.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_.*umma -s 4 -c 2 -o gpurun_out/prof_attn_umma2 -f python scripts/prof_attn.py > gpurun_out/ncu_attn.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_attn.log
