#!/bin/bash
# 1-GPU confirmation call: full -m gpu suite, headline bench (both arms), attention microbench.
set -u
OUT=${GRAFT_REPO_ROOT:-.}/gpurun_out
mkdir -p "$OUT"; : > "$OUT/summary.txt"
timeout 1000 python -m pytest tests -m gpu -x -q -p no:cacheprovider > "$OUT/pytest_gpu.log" 2>&1; echo "pytest rc=$?" | tee -a "$OUT/summary.txt"
tail -n 5 "$OUT/pytest_gpu.log"
timeout 900 python bench.py --steps 10 --warmup 3 > "$OUT/bench_n1.json" 2> "$OUT/bench_n1.err"; echo "bench rc=$?" | tee -a "$OUT/summary.txt"
tail -c 4000 "$OUT/bench_n1.json"; tail -n 5 "$OUT/bench_n1.err"
timeout 400 python tools/bench_kernels.py --only attn > "$OUT/attn_micro.jsonl" 2> "$OUT/attn_micro.err"; echo "attn micro rc=$?" | tee -a "$OUT/summary.txt"
cat "$OUT/attn_micro.jsonl"
cat "$OUT/summary.txt"
