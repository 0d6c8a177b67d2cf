#!/bin/bash
# 4-GPU check of the final tree: TP4 bench (both arms) with the first inference cases
set -u
OUT=${GRAFT_REPO_ROOT:-.}/gpurun_out
mkdir -p "$OUT"
NXD_BENCH_INFERENCE_BUDGET=${NXD_BENCH_INFERENCE_BUDGET:-70} timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 4 --steps 5 --warmup 3 > "$OUT/bench_tp4.log" 2>&1; echo "bench rc=$?"
tail -n 1 "$OUT/bench_tp4.log" > "$OUT/bench_n4.json"
python - <<'PY'
import json
try:
    d = json.load(open("gpurun_out/bench_n4.json"))
    print("TP4 ours", d["ms_per_step"], "library", d["baseline_arm"]["ms_per_step"], "ratio", d["vs_baseline"], "tok/s", d["value"], "fallbacks", d["fallbacks"], "fused_check", d["fused_check"])
    inf = d.get("inference", {})
    if "error" in inf: print("inference error", inf["error"])
    for c in inf.get("cases", []):
        if "ours" in c:
            print(c["model"], c["batch"], "ours ttft %.2f dec %.3f | lib ttft %.2f dec %.3f | ratio %.2f" % (c["ours"]["ttft_ms_p50"], c["ours"]["decode_ms_per_token_p50"], c["library"]["ttft_ms_p50"], c["library"]["decode_ms_per_token_p50"], c["vs_baseline"]["decode_tokens_per_s"]))
        else:
            print(c)
except Exception as e:
    print("bench parse failed", e)
    import subprocess; print(subprocess.run("tail -n 20 gpurun_out/bench_tp4.log", shell=True, capture_output=True, text=True).stdout[-3000:])
PY
