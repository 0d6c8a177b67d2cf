#!/bin/bash
# 2-GPU regression of the decode changes (small-M GEMM routing under TP, PDL in the all-reduce GEMV kernels, tensor-core
# decode attention at TP head counts): TP2 inference tests + TP2 bench.  Tight limits per step.
set -u
OUT=${GRAFT_REPO_ROOT:-.}/gpurun_out
mkdir -p "$OUT"
step() { local name=$1 limit=$2; shift 2; echo "== $name"; timeout "$limit" "$@" > "$OUT/$name.log" 2>&1; local rc=$?; echo "$name rc=$rc  $(tail -n 1 "$OUT/$name.log" | cut -c1-200)" | tee -a "$OUT/summary_2gpu_c.txt"; return $rc; }
: > "$OUT/summary_2gpu_c.txt"
step tp2_inference_tests 240 python -m pytest tests/test_inference_gpu.py -m gpu -x -q -p no:cacheprovider || exit 1
step fused_comm_tests 300 python -m pytest tests/test_fused_comm_gpu.py -m gpu -x -q -p no:cacheprovider
step bench_tp2 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3
tail -n 1 "$OUT/bench_tp2.log" > "$OUT/bench_n2.json"
python - <<'PY'
import json
try:
    d = json.load(open("gpurun_out/bench_n2.json"))
    print("TP2 ours", d["ms_per_step"], "library", d["baseline_arm"]["ms_per_step"], "ratio", d["vs_baseline"], "fallbacks", d["fallbacks"], "fused_check", d["fused_check"])
    for c in d["inference"]["cases"]:
        print(c["model"], c["batch"], "ours ttft %.2f dec %.3f | lib ttft %.2f dec %.3f | ratio %.2f" % (c["ours"]["ttft_ms_p50"], c["ours"]["decode_ms_per_token_p50"], c["library"]["ttft_ms_p50"], c["library"]["decode_ms_per_token_p50"], c["vs_baseline"]["decode_tokens_per_s"]))
except Exception as e:
    print("bench parse failed", e)
PY
