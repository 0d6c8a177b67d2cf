#!/bin/bash
# programmatic dependent launch A/B: decode + inference numerics, then the inference bench with and without the attribute
set -u
OUT=${GRAFT_REPO_ROOT:-.}/gpurun_out
mkdir -p "$OUT"
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_inference_gpu.py -m gpu -x -q -p no:cacheprovider > "$OUT/pdl_test.log" 2>&1; echo "test rc=$?"
tail -n 6 "$OUT/pdl_test.log"
for P in 1 0; do
  for B in 1 16 64; do
    NXD_PDL=$P timeout 200 python tools/bench_inference.py --model llama3.2-1b --batch $B --prompt 2048 --new 64 --runs 10 2>&1 | tail -n 1 | cut -c1-400 | sed "s/^/PDL=$P /"
  done
done
