#!/bin/bash
# one ncu --set full capture of the batched token-generation kernels (after the same program ran clean without ncu)
set -u
OUT=${GRAFT_REPO_ROOT:-.}/gpurun_out
mkdir -p "$OUT"
timeout 120 python tools/profile_target.py decode_batched > "$OUT/plain_decode_batched.log" 2>&1 || { tail -n 5 "$OUT/plain_decode_batched.log"; exit 1; }
timeout 400 ncu --set full --clock-control none --import-source on -k regex:"gemm_smallm_kernel|decode_attn_mma_kernel" \
    --launch-skip 130 -c 10 -f -o "$OUT/decode_batched_prof" python tools/profile_target.py decode_batched > "$OUT/ncu_decode_batched.log" 2>&1; echo "ncu rc=$?"
tail -n 3 "$OUT/ncu_decode_batched.log"
ls -la "$OUT"/decode_batched_prof.ncu-rep
