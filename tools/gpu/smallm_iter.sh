#!/bin/bash
# small-M GEMM iteration: numerics of the kernel family and of the batched decode layer, microbench, inference bench
set -u
OUT=${GRAFT_REPO_ROOT:-.}/gpurun_out
mkdir -p "$OUT"
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_inference_gpu.py -m gpu -x -q -k "smallm or batched_decode or generation_matches or cuda_graphs" -p no:cacheprovider > "$OUT/smallm_test.log" 2>&1; echo "test rc=$?"
tail -n 15 "$OUT/smallm_test.log"
if [ "${SKIP_MICRO:-0}" != "1" ]; then
timeout 300 python tools/bench_kernels.py --only smallm > "$OUT/smallm_micro.jsonl" 2> "$OUT/smallm_micro.err"; echo "micro rc=$?"
python - <<'PY'
import json,os
for l in open(os.path.join(os.environ.get("GRAFT_REPO_ROOT","."),"gpurun_out/smallm_micro.jsonl")):
    r=json.loads(l)
    sp=" ".join(f"s{k}={r[f'ours_s{k}_ms']*1e3:.1f}" for k in (1,2,4,8) if f"ours_s{k}_ms" in r)
    print(f"{r['case']:12s} B={r['B']:3d} ours {r['ours_ms']*1e3:6.1f} us ({sp}) cublas {r['cublas_ms']*1e3:6.1f} prev {r['prev_ms']*1e3:6.1f} roof {r['roofline_ms']*1e3:5.1f} err {r['max_err']:.3f}")
PY
tail -n 5 "$OUT/smallm_micro.err"
fi
timeout 600 python bench.py --steps 3 --warmup 3 > "$OUT/bench_smallm.json" 2> "$OUT/bench_smallm.err"; echo "bench rc=$?"
python - <<'PY'
import json,os
d=json.loads(open(os.path.join(os.environ.get("GRAFT_REPO_ROOT","."),"gpurun_out/bench_smallm.json")).read().strip().splitlines()[-1])
print("train ms", d["ms_per_step"], "lib", d["baseline_arm"]["ms_per_step"], "ratio", d["vs_baseline"])
for c in d["inference"]["cases"]:
    print(c["model"], "b", c["batch"], "ours ttft %.2f dec %.3f | lib ttft %.2f dec %.3f | ratio dec %.2f" % (c["ours"]["ttft_ms_p50"], c["ours"]["decode_ms_per_token_p50"], c["library"]["ttft_ms_p50"], c["library"]["decode_ms_per_token_p50"], c["vs_baseline"]["decode_tokens_per_s"]))
PY
tail -n 5 "$OUT/bench_smallm.err"
