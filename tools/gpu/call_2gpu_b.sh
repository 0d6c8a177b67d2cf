#!/bin/bash
# 2-GPU session B: ZeRO-1 with DP = 2 on a measured path, PP = 2 over NCCL p2p at the 3B size, Mixtral EP = 2 training.
set -u
OUT=${GRAFT_REPO_ROOT:-.}/gpurun_out
mkdir -p "$OUT"
step() { local name=$1 limit=$2; shift 2; echo "== $name"; timeout "$limit" "$@" > "$OUT/$name.log" 2>&1; local rc=$?; echo "$name rc=$rc  $(tail -n 1 "$OUT/$name.log" | cut -c1-300)" | tee -a "$OUT/summary_2gpu_b.txt"; }
: > "$OUT/summary_2gpu_b.txt"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
step bench_dp2 600 $TR --master-port 29521 bench.py --gpus 2 --tp 1 --pp 1 --steps 5 --warmup 3 --no-inference
step bench_pp2 600 $TR --master-port 29522 bench.py --gpus 2 --tp 1 --pp 2 --microbatches 4 --steps 5 --warmup 3 --no-inference
step ep2_train 400 $TR --master-port 29523 examples/training/mixtral_pretrain.py --layers 2 --expert_parallel_size 2 --seq_len 2048 --max_steps 5
for f in bench_dp2 bench_pp2; do tail -n 1 "$OUT/$f.log" > "$OUT/$f.json"; done
python - <<'PY'
import json
for f in ("bench_dp2", "bench_pp2"):
    try:
        d = json.load(open(f"gpurun_out/{f}.json"))
        print(f, d["config"]["parallelism"], "ours", round(d["ms_per_step"], 1), "ms  library", round(d["baseline_arm"]["ms_per_step"], 1), "ms  ratio", round(d["vs_baseline"], 3), "tok/s", round(d["value"]), "fallbacks", d["fallbacks"])
    except Exception as e:
        print(f, "parse failed", e)
PY
tail -n 5 "$OUT/ep2_train.log"
