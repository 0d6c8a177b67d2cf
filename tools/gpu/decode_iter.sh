#!/bin/bash
set -u
OUT=${GRAFT_REPO_ROOT:-.}/gpurun_out
mkdir -p "$OUT"
timeout 180 python -m pytest tests/test_kernels_gpu.py tests/test_inference_gpu.py -m gpu -x -q -p no:cacheprovider -k "${NXD_TEST_FILTER:-decode or inference or skinny or sampling}" > "$OUT/decode_tests.log" 2>&1; echo "decode tests rc=$?"
tail -n 8 "$OUT/decode_tests.log"
for B in 1 16 64; do
  timeout 120 python tools/bench_inference.py --model llama3.2-1b --batch $B --prompt 2048 --new 64 --runs 5 2>&1 | tail -n 1
done
timeout 120 python tools/bench_inference.py --model llama3.2-1b --batch 16 --prompt 2048 --new 32 --runs 2 --profile "$OUT/decode_kernels_tp1_b16.txt" > /dev/null 2>&1; head -n 14 "$OUT/decode_kernels_tp1_b16.txt"
timeout 120 python tools/bench_inference.py --model llama3.2-1b --batch 1 --prompt 2048 --new 32 --runs 2 --profile "$OUT/decode_kernels_tp1_b1.txt" > /dev/null 2>&1; head -n 22 "$OUT/decode_kernels_tp1_b1.txt"
