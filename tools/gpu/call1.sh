#!/bin/bash
# Round-2 first GPU call (1 GPU): full -m gpu suite, headline bench with the in-job library arm, attention microbench,
# first hardware run of the int8 skinny GEMM, ncu capture of the attention kernels.
set -u
OUT=${GRAFT_REPO_ROOT:-.}/gpurun_out
mkdir -p "$OUT"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > "$OUT/smi_start.csv" 2>&1
timeout 900 python -m pytest tests -m gpu -x -q -p no:cacheprovider > "$OUT/pytest_gpu.log" 2>&1; echo "pytest rc=$?" | tee -a "$OUT/summary.txt"
tail -n 5 "$OUT/pytest_gpu.log"
NXD_UNVERIFIED=1 timeout 300 python -m pytest tests/test_kernels_gpu.py -q -m gpu -k skinny_gemm_i8 -p no:cacheprovider > "$OUT/i8.log" 2>&1; echo "i8 rc=$?" | tee -a "$OUT/summary.txt"
tail -n 3 "$OUT/i8.log"
timeout 900 python bench.py --steps 10 --warmup 3 > "$OUT/bench_n1.json" 2> "$OUT/bench_n1.err"; echo "bench rc=$?" | tee -a "$OUT/summary.txt"
tail -c 3000 "$OUT/bench_n1.json"; tail -n 5 "$OUT/bench_n1.err"
timeout 600 python tools/bench_kernels.py --only attn > "$OUT/attn_micro.jsonl" 2> "$OUT/attn_micro.err"; echo "attn micro rc=$?" | tee -a "$OUT/summary.txt"
cat "$OUT/attn_micro.jsonl"
timeout 300 python tools/profile_target.py attn > "$OUT/plain_attn.log" 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fa_.*_tc_kernel -c 4 -f -o "$OUT/attn_prof" \
    python tools/profile_target.py attn > "$OUT/ncu_attn.log" 2>&1; echo "ncu rc=$?" | tee -a "$OUT/summary.txt"
tail -n 5 "$OUT/ncu_attn.log"
cat "$OUT/summary.txt"
