#!/bin/bash
set -u
OUT=${GRAFT_REPO_ROOT:-.}/gpurun_out
mkdir -p "$OUT"
timeout 300 python tools/profile_target.py attn > "$OUT/plain_attn.log" 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fa_.*_tc_kernel -c 2 -f -o "$OUT/attn_prof" \
    python tools/profile_target.py attn > "$OUT/ncu_attn.log" 2>&1; echo "ncu rc=$?"
tail -n 3 "$OUT/ncu_attn.log"
