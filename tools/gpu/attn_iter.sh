#!/bin/bash
# attention kernel iteration: numerics tests then the microbenchmark
set -u
OUT=${GRAFT_REPO_ROOT:-.}/gpurun_out
mkdir -p "$OUT"
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -p no:cacheprovider -k "${NXD_TEST_FILTER:-flash_attention}" > "$OUT/attn_tests.log" 2>&1; echo "attn tests rc=$?"
tail -n 15 "$OUT/attn_tests.log"
timeout 600 python tools/bench_kernels.py --only attn > "$OUT/attn_micro.jsonl" 2> "$OUT/attn_micro.err"; echo "attn micro rc=$?"
python - <<'PY'
import json
for l in open("gpurun_out/attn_micro.jsonl"):
    r = json.loads(l)
    print(f"{r['case']:18s} fwd {r['nxd_fwd_ms']:.3f} ms {r['nxd_fwd_tflops']:6.0f} TF (sdpa {r['sdpa_fwd_tflops']:6.0f})  bwd {r['nxd_bwd_ms']:.3f} ms {r['nxd_bwd_tflops']:6.0f} TF (sdpa {r['sdpa_bwd_tflops']:6.0f})")
PY
