#!/bin/bash
set -u
OUT=${GRAFT_REPO_ROOT:-.}/gpurun_out
mkdir -p "$OUT"
# tight limits: a hang of a new kernel must cost seconds, not the call's whole budget
timeout 60 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "gemm_smallm and 16-2048-3072" -p no:cacheprovider > "$OUT/smallm_first.log" 2>&1 || { echo "first case failed / hung"; tail -n 5 "$OUT/smallm_first.log"; exit 1; }
timeout 150 python -m pytest tests/test_kernels_gpu.py tests/test_inference_gpu.py -m gpu -x -q -k "smallm or batched_decode" -p no:cacheprovider > "$OUT/smallm_test.log" 2>&1; rc=$?; echo "test rc=$rc"
tail -n 8 "$OUT/smallm_test.log"
[ $rc -eq 0 ] || exit 1
NXD_B200_LIB=$GRAFT_REPO_ROOT/build/libnxd_trace.so timeout 60 python tools/gpu/smallm_trace.py
NXD_SMALLM_B=16,64 timeout 120 python tools/bench_kernels.py --only smallm > "$OUT/smallm_micro.jsonl" 2> "$OUT/smallm_micro.err"; echo "micro rc=$?"
python - <<'PY'
import json,os
for l in open(os.path.join(os.environ.get("GRAFT_REPO_ROOT","."),"gpurun_out/smallm_micro.jsonl")):
    r=json.loads(l)
    sp=" ".join(f"s{k}={r[f'ours_s{k}_ms']*1e3:.1f}" for k in (1,2,4,8) if f"ours_s{k}_ms" in r)
    print(f"{r['case']:12s} B={r['B']:3d} ours {r['ours_ms']*1e3:6.1f} us ({sp}) cublas {r['cublas_ms']*1e3:6.1f} prev {r['prev_ms']*1e3:6.1f} roof {r['roofline_ms']*1e3:5.1f} err {r['max_err']:.3f}")
PY
tail -n 5 "$OUT/smallm_micro.err"
