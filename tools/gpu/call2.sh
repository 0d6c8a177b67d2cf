#!/bin/bash
# kernel timelines of both arms (CUPTI through torch.profiler; never a bench value) + compute-sanitizer memcheck
set -u
OUT=${GRAFT_REPO_ROOT:-.}/gpurun_out
mkdir -p "$OUT"
timeout 600 python bench.py --steps 2 --warmup 3 --no-inference --profile "$OUT/timeline_tp1_native.txt" > "$OUT/prof_native.log" 2>&1; echo "profile native rc=$?"
timeout 600 python bench.py --impl baseline --steps 2 --warmup 3 --no-inference --profile "$OUT/timeline_tp1_library.txt" > "$OUT/prof_lib.log" 2>&1; echo "profile library rc=$?"
head -n 40 "$OUT/timeline_tp1_native.txt"; head -n 40 "$OUT/timeline_tp1_library.txt"
SANITIZER_TIMEOUT=900 timeout 1000 bash tools/run_sanitizers.sh memcheck "rmsnorm or swiglu or rope or skinny or sampling or cross_entropy or embedding or adamw"; echo "memcheck rc=$?"
