#!/bin/bash
# full -m gpu suite (tight limit) + headline bench with both arms; prints a digest
set -u
OUT=${GRAFT_REPO_ROOT:-.}/gpurun_out
mkdir -p "$OUT"
timeout 400 python -m pytest tests -m gpu -x -q -p no:cacheprovider > "$OUT/pytest_gpu.log" 2>&1; echo "pytest rc=$?"
tail -n 3 "$OUT/pytest_gpu.log"
timeout 500 python bench.py --steps ${STEPS:-10} --warmup 3 > "$OUT/bench_n1.json" 2> "$OUT/bench_n1.err"; echo "bench rc=$?"
python - <<'PY'
import json,os
d=json.loads(open(os.path.join(os.environ.get("GRAFT_REPO_ROOT","."),"gpurun_out/bench_n1.json")).read().strip().splitlines()[-1])
print("train ms", round(d["ms_per_step"],2), "lib", round(d["baseline_arm"]["ms_per_step"],2), "ratio", round(d["vs_baseline"],3), "clocks", d["clocks"]["sm_mhz"], d["baseline_arm"]["clocks"]["sm_mhz"], "launches", d["gpu_launches"], "fallbacks", d["fallbacks"])
for c in d["inference"]["cases"]:
    print(c["model"], "b", c["batch"], "ours ttft %.2f dec %.3f | lib ttft %.2f dec %.3f | ratio ttft %.2f dec %.2f" % (c["ours"]["ttft_ms_p50"], c["ours"]["decode_ms_per_token_p50"], c["library"]["ttft_ms_p50"], c["library"]["decode_ms_per_token_p50"], c["vs_baseline"]["ttft"], c["vs_baseline"]["decode_tokens_per_s"]))
PY
tail -n 3 "$OUT/bench_n1.err"
