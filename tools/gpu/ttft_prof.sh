#!/bin/bash
set -u
OUT=${GRAFT_REPO_ROOT:-.}/gpurun_out
mkdir -p "$OUT"
for impl in ours baseline; do
  timeout 150 python tools/bench_inference.py --model llama3.2-3b --impl $impl --batch 1 --prompt 2048 --new 2 --runs 2 --profile "$OUT/prefill_kernels_3b_b1_$impl.txt" > /dev/null 2>&1; head -n 16 "$OUT/prefill_kernels_3b_b1_$impl.txt" | cut -c1-140
done
