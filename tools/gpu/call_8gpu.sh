#!/bin/bash
# 8-GPU session: fused-collective numerics at world 8, headline TP8 bench (+ inference) with the in-job library arm,
# NVLS all-gather variant, BASELINE config #3 (TP4 x PP2 + activation checkpointing), fused kernel roofline table.
set -u
OUT=${GRAFT_REPO_ROOT:-.}/gpurun_out
mkdir -p "$OUT"
step() { local name=$1 limit=$2; shift 2; echo "== $name"; timeout "$limit" "$@" > "$OUT/$name.log" 2>&1; local rc=$?; echo "$name rc=$rc  $(tail -n 1 "$OUT/$name.log" | cut -c1-200)" | tee -a "$OUT/summary_8gpu.txt"; }
: > "$OUT/summary_8gpu.txt"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
step tests_world8 400 python -m pytest tests/test_fused_comm_gpu.py -m gpu -x -q -p no:cacheprovider -k "8 and (ag_gemm or gemm_rs or gemm_ar or embedding)"
NXD_STRESS=1 NXD_STRESS_WORLD=8 step flag_stress_8gpu 300 python -m pytest tests/test_fused_comm_gpu.py -q -m gpu -k flag_protocol_stress -p no:cacheprovider
step bench_tp8 900 $TR --master-port 29531 bench.py --gpus 8 --steps 10 --warmup 3
NXD_NVLS=1 step bench_tp8_nvls 400 $TR --master-port 29532 bench.py --gpus 8 --steps 10 --warmup 3 --no-inference --no-baseline --no-e2e
step bench_tp4pp2 600 $TR --master-port 29533 bench.py --gpus 8 --tp 4 --pp 2 --microbatches 4 --act-ckpt --steps 5 --warmup 3 --no-inference --no-e2e
step fused_tp8 300 $TR --master-port 29534 tools/bench_fused.py
NXD_AG_BLOCK_N=256 step fused_tp8_bn256 300 $TR --master-port 29535 tools/bench_fused.py
for f in bench_tp8 bench_tp8_nvls bench_tp4pp2; do tail -n 1 "$OUT/$f.log" > "$OUT/$f.json"; done
python - <<'PY'
import json
for f in ("bench_tp8", "bench_tp8_nvls", "bench_tp4pp2"):
    try:
        d = json.load(open(f"gpurun_out/{f}.json"))
        b = d.get("baseline_arm") or {}
        print(f, d["config"]["parallelism"], "ours", round(d["ms_per_step"], 2), "ms  library", b.get("ms_per_step"), "ratio", d["vs_baseline"], "tok/s", round(d["value"]), "fallbacks", d["fallbacks"], "fused_check", d.get("fused_check"))
        for c in (d.get("inference") or {}).get("cases", []):
            print("   ", c["model"], c["batch"], c.get("ours"), c.get("vs_baseline"))
    except Exception as e:
        print(f, "parse failed", e)
PY
grep -h "case" "$OUT/fused_tp8.log" | cut -c1-400
echo "-- bn256"; grep -h "case" "$OUT/fused_tp8_bn256.log" | cut -c1-400
cat "$OUT/summary_8gpu.txt"
