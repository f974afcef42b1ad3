#!/bin/bash
mkdir -p gpurun_out
timeout 180 python scripts/dev_attn.py 64 > gpurun_out/attn.log 2>&1; echo "attn rc=$?"; grep -E "rel|time|identical" gpurun_out/attn.log | head -6
for d in 0 32 64 96 112 15; do S4G_ATTN_DBG=$d timeout 60 python scripts/dev_attn_time.py 2>&1 | grep "dbg.* b "; done | tee gpurun_out/attn_dbg5.log
