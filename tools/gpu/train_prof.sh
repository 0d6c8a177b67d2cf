#!/bin/bash
set -u
OUT=${GRAFT_REPO_ROOT:-.}/gpurun_out
mkdir -p "$OUT"
timeout 200 python bench.py --impl ours --no-inference --profile "$OUT/timeline_tp1_native.txt" > /dev/null 2> "$OUT/prof_ours.err"; echo "rc=$?"
timeout 200 python bench.py --impl baseline --no-inference --profile "$OUT/timeline_tp1_library.txt" > /dev/null 2> "$OUT/prof_lib.err"; echo "rc=$?"
head -n 32 "$OUT/timeline_tp1_native.txt" | cut -c1-150
head -n 32 "$OUT/timeline_tp1_library.txt" | cut -c1-150
