#!/bin/bash
set -u
OUT=${GRAFT_REPO_ROOT:-.}/gpurun_out
mkdir -p "$OUT"
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_inference_gpu.py -m gpu -x -q -p no:cacheprovider -k "decode or inference or generation or batched or cuda_graphs" > "$OUT/decattn_test.log" 2>&1; echo "test rc=$?"
tail -n 12 "$OUT/decattn_test.log"
for A in mma fma; do
  for B in 1 16 64; do
    NXD_DECODE_ATTN=$A timeout 200 python tools/bench_inference.py --model llama3.2-1b --batch $B --prompt 2048 --new 64 --runs 10 2>&1 | tail -n 1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$A', 'b', d['batch'], 'ttft', round(d['ttft_ms_p50'],2), 'ms/token', round(d['decode_ms_per_token_p50'],4), 'tok/s', round(d['decode_tokens_per_s']))"
  done
done
NXD_DECODE_ATTN=mma timeout 200 python tools/bench_inference.py --model llama3.2-3b --batch 16 --prompt 2048 --new 64 --runs 5 2>&1 | tail -n 1 | cut -c1-330
timeout 120 python tools/bench_inference.py --model llama3.2-1b --batch 64 --prompt 2048 --new 32 --runs 2 --profile "$OUT/decode_kernels_tp1_b64.txt" > /dev/null 2>&1; head -n 12 "$OUT/decode_kernels_tp1_b64.txt" | cut -c1-130
