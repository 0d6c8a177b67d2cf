#!/bin/bash
set -u
OUT=${GRAFT_REPO_ROOT:-.}/gpurun_out
mkdir -p "$OUT"
for B in 16 64; do
  timeout 120 python tools/bench_inference.py --model llama3.2-1b --batch $B --prompt 2048 --new 32 --runs 2 --profile "$OUT/decode_kernels_tp1_b$B.txt" > /dev/null 2>&1; head -n 16 "$OUT/decode_kernels_tp1_b$B.txt"
done
timeout 120 python tools/bench_inference.py --model llama3.2-1b --impl baseline --batch 64 --prompt 2048 --new 32 --runs 2 --profile "$OUT/decode_kernels_tp1_b64_library.txt" > /dev/null 2>&1; head -n 16 "$OUT/decode_kernels_tp1_b64_library.txt"
