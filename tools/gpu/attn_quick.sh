#!/bin/bash
set -u
OUT=${GRAFT_REPO_ROOT:-.}/gpurun_out
mkdir -p "$OUT"
timeout 200 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "flash_attention" -p no:cacheprovider > "$OUT/attn_test.log" 2>&1; echo "test rc=$?"
tail -n 4 "$OUT/attn_test.log"
timeout 200 python tools/bench_kernels.py --only attn > "$OUT/attn_micro.jsonl" 2> "$OUT/attn_micro.err"; echo "micro rc=$?"
python - <<'PY'
import json,os
for l in open(os.path.join(os.environ.get("GRAFT_REPO_ROOT","."),"gpurun_out/attn_micro.jsonl")):
    r=json.loads(l)
    print(r["case"], "fwd %.3f (%.0f TF) bwd %.3f (%.0f TF) | sdpa fwd %.3f bwd %.3f" % (r["nxd_fwd_ms"], r["nxd_fwd_tflops"], r["nxd_bwd_ms"], r["nxd_bwd_tflops"], r["sdpa_fwd_ms"], r["sdpa_bwd_ms"]))
PY
