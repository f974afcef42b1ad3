#!/bin/bash
mkdir -p gpurun_out
timeout 180 python scripts/dev_attn.py 3 64 > gpurun_out/attn.log 2>&1; echo "attn rc=$?"; grep -E "rel|time|identical|Error|error" gpurun_out/attn.log | head -30
for d in 0 32 64 96; do S4G_ATTN_DBG=$d timeout 120 python scripts/dev_attn_time.py 12 64 2>&1 | grep "dbg"; done | tee gpurun_out/attn_dbg9.log
