#!/bin/bash
# trajectory test, long-sequence runs with peak memory, ncu captures of the hot kernels (one ncu invocation)
set -u
OUT=${GRAFT_REPO_ROOT:-.}/gpurun_out
mkdir -p "$OUT"
timeout 400 python -m pytest tests/test_trajectory_gpu.py -m gpu -x -q -p no:cacheprovider > "$OUT/trajectory.log" 2>&1; echo "trajectory rc=$?"; tail -n 4 "$OUT/trajectory.log"
for S in 16384 32768; do
  timeout 400 python bench.py --seq-len $S --steps 3 --warmup 3 --no-inference --no-e2e > "$OUT/bench_seq$S.json" 2> "$OUT/bench_seq$S.err"; echo "seq $S rc=$?"
  python -c "
import json; d=json.load(open('$OUT/bench_seq$S.json')); b=d['baseline_arm']
print('S=$S ours', round(d['ms_per_step'],1), 'ms', round(d['value']), 'tok/s peak', round(d['peak_mem_gb'],1), 'GB | library', round(b['ms_per_step'],1), 'ms peak', round(b['peak_mem_gb'],1), 'GB ratio', round(d['vs_baseline'],3))" || tail -n 3 "$OUT/bench_seq$S.err"
done
timeout 300 python tools/profile_target.py all > "$OUT/plain_all.log" 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'gemm_pair_kernel|fa_fwd_tc|fa_bwd_tc_kernel|rmsnorm_fwd|rmsnorm_bwd_kernel|swiglu_fwd|swiglu_bwd|adamw_kernel|ce_stats|decode_attn_split|decode_qkv|decode_gate_up|skinny_splitk' -c 16 -f -o "$OUT/hot_kernels" \
    python tools/profile_target.py all > "$OUT/ncu_all.log" 2>&1; echo "ncu rc=$?"; tail -n 3 "$OUT/ncu_all.log"
