#!/bin/bash
# round 2, call B: tcgen05 attention bring-up (its own timeout: a hung kernel must not take the box down) + conditioning sweeps
set -x
mkdir -p gpurun_out
timeout 120 python scripts/dev_attn.py 1 > gpurun_out/attn_b1.log 2>&1; echo "attn B=1 rc=$?"; cat gpurun_out/attn_b1.log | tail -12
timeout 180 python scripts/dev_attn.py 3 64 > gpurun_out/attn.log 2>&1; echo "attn rc=$?"; cat gpurun_out/attn.log | tail -30
S4G_ATTN=mma timeout 120 python scripts/dev_attn.py 64 > gpurun_out/attn_mma.log 2>&1; grep time gpurun_out/attn_mma.log
for d in seg depth; do timeout 400 python tests/tools/dev_wellcond.py $d 4 10 30 > gpurun_out/wellcond_$d.log 2>&1; grep -v "^   l2 " gpurun_out/wellcond_$d.log; done
timeout 300 python tests/tools/dev_wellcond.py seg 32 30 > gpurun_out/wellcond_seg32.log 2>&1; grep -v "^   l2 " gpurun_out/wellcond_seg32.log
