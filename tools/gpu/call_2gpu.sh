#!/bin/bash
# 2-GPU session: multi-GPU numerics (fused collectives, TP2 inference, PP2 over NCCL p2p), flag-protocol stress,
# the NVLS all-gather variant, and the TP2 bench with the in-job library arm.
set -u
OUT=${GRAFT_REPO_ROOT:-.}/gpurun_out
mkdir -p "$OUT"
step() { local name=$1 limit=$2; shift 2; echo "== $name"; timeout "$limit" "$@" > "$OUT/$name.log" 2>&1; local rc=$?; echo "$name rc=$rc  $(tail -n 1 "$OUT/$name.log" | cut -c1-200)" | tee -a "$OUT/summary_2gpu.txt"; }
: > "$OUT/summary_2gpu.txt"
step multi_gpu_tests 600 python -m pytest tests/test_fused_comm_gpu.py tests/test_inference_gpu.py tests/test_pipeline_gpu.py -m gpu -x -q -p no:cacheprovider
NXD_STRESS=1 step flag_stress_2gpu 400 python -m pytest tests/test_fused_comm_gpu.py -q -m gpu -k flag_protocol_stress -p no:cacheprovider
NXD_NVLS=1 step nvls_ag_gemm 300 python -m pytest tests/test_fused_comm_gpu.py -q -m gpu -k ag_gemm -p no:cacheprovider
step bench_tp2 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3
tail -n 1 "$OUT/bench_tp2.log" > "$OUT/bench_n2.json"
python - <<'PY'
import json
try:
    d = json.load(open("gpurun_out/bench_n2.json"))
    print("TP2 ours", d["ms_per_step"], "library", d["baseline_arm"]["ms_per_step"], "ratio", d["vs_baseline"], "fallbacks", d["fallbacks"], "fused_check", d["fused_check"])
    for c in d["inference"]["cases"]:
        print(c["model"], c["batch"], c.get("ours"), c.get("vs_baseline"))
except Exception as e:
    print("bench parse failed", e)
PY
cat "$OUT/summary_2gpu.txt"
