#!/usr/bin/env python
"""Micro-benchmarks of the hand-written kernels at Llama-3.2 shapes against the library baselines
(cuBLAS via torch.matmul, PyTorch SDPA) and the measured roofline (MEASURED_PEAKS.json).

CUDA events around ITERS back-to-back launches after warm-up; an L2 flush (write of a 512 MB buffer) precedes every
timed launch for the bandwidth-bound ops, GEMM/attention operands are cycled through > L2-sized pools.
Writes one JSON line per case to stdout and (optionally) a markdown table with --md."""
from __future__ import annotations

import argparse
import json
import math
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from neuronx_distributed_llama3_2_b200.ops import _lib  # noqa: E402
from neuronx_distributed_llama3_2_b200.ops import attention as A  # noqa: E402
from neuronx_distributed_llama3_2_b200.ops import backend  # noqa: E402
from neuronx_distributed_llama3_2_b200.ops.gemm import gemm  # noqa: E402

PEAKS = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(
    os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"bf16_tflops": 1590.0, "hbm_gbs": 6650.0}


def time_ms(fn, iters=20, warmup=5, flush=None):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    if flush is None:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters
    tot = 0.0
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / iters


def bench_gemm(rows):
    shapes = [  # (name, layout, M, N, K)
        ("3B qkv fwd tp1", "nt", 8192, 5120, 3072), ("3B gate_up fwd tp1", "nt", 8192, 16384, 3072),
        ("3B down fwd tp1", "nt", 8192, 3072, 8192), ("3B o fwd tp1", "nt", 8192, 3072, 3072),
        ("3B gate_up dgrad tp1", "nn", 8192, 3072, 16384), ("3B gate_up wgrad tp1", "tn", 16384, 3072, 8192),
        ("3B down wgrad tp1", "tn", 3072, 8192, 8192), ("3B lm_head fwd tp1", "nt", 8192, 128256, 3072),
        ("3B gate_up fwd tp8", "nt", 8192, 2048, 3072), ("3B down fwd tp8", "nt", 8192, 3072, 1024),
        ("3B qkv fwd tp8", "nt", 8192, 640, 3072), ("1B gate_up fwd tp8", "nt", 8192, 2048, 2048),
        ("square 8192", "nt", 8192, 8192, 8192), ("square 4096", "nt", 4096, 4096, 4096),
        ("3B qkv wgrad tp1", "tn", 5120, 3072, 8192), ("3B o wgrad tp1", "tn", 3072, 3072, 8192),
        ("3B qkv dgrad tp1", "nn", 8192, 3072, 5120), ("3B o dgrad tp1", "nn", 8192, 3072, 3072),
        ("3B down dgrad tp1", "nn", 8192, 8192, 3072),
        ("3B qkv prefill2048", "nt", 2048, 5120, 3072), ("3B o prefill2048", "nt", 2048, 3072, 3072),
        ("3B gate_up prefill2048", "nt", 2048, 16384, 3072), ("3B down prefill2048", "nt", 2048, 3072, 8192),
        ("1B qkv prefill2048", "nt", 2048, 3072, 2048), ("1B o prefill2048", "nt", 2048, 2048, 2048),
        ("1B gate_up prefill2048", "nt", 2048, 16384, 2048), ("1B down prefill2048", "nt", 2048, 2048, 8192),
        ("1B qkv prefill b16", "nt", 32768, 3072, 2048), ("1B down prefill b16", "nt", 32768, 2048, 8192),
    ]
    only = os.environ.get("NXD_GEMM_CASES", "")
    if only:
        shapes = [s for s in shapes if any(o in s[0] for o in only.split(","))]
    for name, layout, M, N, K in shapes:
        pool = max(2, int(math.ceil(300e6 / ((M * K + N * K) * 2))))
        pool = min(pool, 8)
        if layout == "nt":
            As = [torch.randn(M, K, device="cuda").bfloat16() for _ in range(pool)]
            Bs = [torch.randn(N, K, device="cuda").bfloat16() for _ in range(pool)]
            ref = lambda a, b: torch.matmul(a, b.t())
        elif layout == "nn":
            As = [torch.randn(M, K, device="cuda").bfloat16() for _ in range(pool)]
            Bs = [torch.randn(K, N, device="cuda").bfloat16() for _ in range(pool)]
            ref = lambda a, b: torch.matmul(a, b)
        else:
            As = [torch.randn(K, M, device="cuda").bfloat16() for _ in range(pool)]
            Bs = [torch.randn(K, N, device="cuda").bfloat16() for _ in range(pool)]
            ref = lambda a, b: torch.matmul(a.t(), b)
        out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        i = [0]

        def ours(bn=0):
            i[0] += 1
            gemm(As[i[0] % pool], Bs[i[0] % pool], layout, out=out, block_n=bn)

        def lib():
            i[0] += 1
            ref(As[i[0] % pool], Bs[i[0] % pool])

        fl = 2.0 * M * N * K
        t_lib = time_ms(lib)
        res = {"op": "gemm", "case": name, "layout": layout, "M": M, "N": N, "K": K,
               "cublas_ms": t_lib, "cublas_tflops": fl / t_lib / 1e9}
        best = None
        for bn in (128, 256):
            t = time_ms(lambda: ours(bn))
            res[f"ours_bn{bn}_ms"] = t
            best = t if best is None else min(best, t)
        t_auto = time_ms(lambda: ours(0))
        res.update({"ours_ms": t_auto, "ours_tflops": fl / t_auto / 1e9, "ours_best_ms": best,
                    "frac_of_measured_peak": fl / t_auto / 1e9 / PEAKS["bf16_tflops"], "vs_cublas": t_lib / t_auto})
        rows.append(res)
        print(json.dumps(res), flush=True)
        del As, Bs, out


def bench_attention(rows):
    for name, B, Hq, Hkv, S, D in [("3B tp1 S8192", 1, 24, 8, 8192, 128), ("3B tp8 S8192", 1, 3, 1, 8192, 128),
                                   ("1B tp1 S8192", 1, 32, 8, 8192, 64), ("3B tp1 S4096 B2", 2, 24, 8, 4096, 128)]:
        q = torch.randn(B, Hq, S, D, device="cuda").bfloat16().requires_grad_()
        k = torch.randn(B, Hkv, S, D, device="cuda").bfloat16().requires_grad_()
        v = torch.randn(B, Hkv, S, D, device="cuda").bfloat16().requires_grad_()
        do = torch.randn(B, Hq, S, D, device="cuda").bfloat16()
        fl_f = 4.0 * B * Hq * S * S * D * 0.5
        res = {"op": "flash_attn", "case": name, "B": B, "Hq": Hq, "Hkv": Hkv, "S": S, "D": D}
        for be in ("nxd", "sdpa"):
            backend.set_backends(attn=be)
            o = A.flash_attn_func(q, k, v, causal=True)
            tf = time_ms(lambda: A.flash_attn_func(q, k, v, causal=True))

            def fb():
                oo = A.flash_attn_func(q, k, v, causal=True)
                oo.backward(do)
                q.grad = k.grad = v.grad = None

            tfb = time_ms(fb, iters=10, warmup=3)
            res[f"{be}_fwd_ms"] = tf
            res[f"{be}_fwd_tflops"] = fl_f / tf / 1e9
            res[f"{be}_bwd_ms"] = tfb - tf
            res[f"{be}_bwd_tflops"] = 2.5 * fl_f / max(tfb - tf, 1e-6) / 1e9
            del o
        backend.set_backends(attn="nxd")
        res["fwd_frac_of_measured_peak"] = res["nxd_fwd_tflops"] / PEAKS["bf16_tflops"]
        rows.append(res)
        print(json.dumps(res), flush=True)


def bench_elementwise(rows):
    from neuronx_distributed_llama3_2_b200.ops import norm, activation, optim, cross_entropy as ce

    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    T, H, I = 8192, 3072, 8192
    x = torch.randn(T, H, device="cuda").bfloat16()
    r = torch.randn(T, H, device="cuda").bfloat16()
    w = torch.ones(H, device="cuda").bfloat16()
    cases = []
    cases.append(("rmsnorm_fwd 8192x3072", lambda: norm.rms_norm(x, w, 1e-5), 2 * T * H * 2))
    cases.append(("add_rmsnorm_fwd 8192x3072", lambda: norm.fused_add_rms_norm(x, r, w, 1e-5), 4 * T * H * 2))
    gu = torch.randn(T, 2 * I, device="cuda").bfloat16()
    cases.append(("swiglu_fwd 8192x8192", lambda: activation.swiglu(gu), 3 * T * I * 2))
    n = 400_000_000
    master = torch.randn(n, device="cuda"); m = torch.zeros(n, device="cuda"); v2 = torch.zeros(n, device="cuda")
    g = torch.randn(n, device="cuda").bfloat16(); po = torch.empty(n, device="cuda", dtype=torch.bfloat16)
    cases.append(("adamw 400M", lambda: optim.adamw_step_(master, m, v2, g, po, 1e-3, 0.9, 0.95, 1e-8, 0.1, 1),
                  n * (12 + 2 + 12 + 2)))
    s = torch.zeros(1, device="cuda")
    cases.append(("sumsq bf16 400M", lambda: optim.sumsq_(g, s), n * 2))
    V = 128256
    logits = torch.randn(4096, V, device="cuda").bfloat16()
    tgt = torch.randint(0, V, (4096,), device="cuda")
    cases.append(("ce_stats 4096x128256", lambda: ce.ce_local_stats(logits, tgt, 0), 4096 * V * 2))
    for name, fn, bytes_ in cases:
        t = time_ms(fn, iters=10, warmup=3, flush=flush)
        res = {"op": "bw", "case": name, "ms": t, "gbs": bytes_ / t / 1e6,
               "frac_of_measured_hbm": bytes_ / t / 1e6 / PEAKS["hbm_gbs"]}
        rows.append(res)
        print(json.dumps(res), flush=True)


def graph_time_ms(fn, calls=40, replays=5):
    """Device time per call of ``fn`` with the host out of the picture: ``calls`` invocations captured into one CUDA
    graph (as the token-generation loop runs them), replayed ``replays`` times between two events."""
    st = torch.cuda.Stream()
    st.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(st):
        for _ in range(3):
            fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=st):
        for _ in range(calls):
            fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(replays):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (replays * calls)


def bench_smallm(rows):
    """Token-generation GEMMs (9..128 rows) of Llama-3.2-1B / 3B: the weight-streaming tcgen05 kernel against cuBLAS, the
    previous path (mma.sync skinny kernel up to 16 rows, 256-row tcgen05 tiles above) and the HBM roofline.  Weights are
    cycled through a pool > 2 x L2 so that every call streams them from HBM, as consecutive layers do; calls are timed
    inside a CUDA graph (the generation loop is one), so host launch cost is excluded for every contender."""
    from neuronx_distributed_llama3_2_b200.ops.fused_comm import skinny_gemm
    from neuronx_distributed_llama3_2_b200.ops.gemm import gemm_smallm

    shapes = [("1B qkv", 3072, 2048), ("1B o", 2048, 2048), ("1B gate_up", 16384, 2048), ("1B down", 2048, 8192),
              ("1B lm_head", 128256, 2048), ("3B gate_up", 16384, 3072), ("3B down", 3072, 8192)]
    only_b = [int(v) for v in os.environ.get("NXD_SMALLM_B", "16,32,64,128").split(",")]
    for name, N, K in shapes:
        copies = max(2, min(64, (320 << 20) // (N * K * 2) + 1))
        ws = [torch.randn(N, K, device="cuda").bfloat16() * 0.02 for _ in range(copies)]
        for B in only_b:
            x = torch.randn(B, K, device="cuda").bfloat16()
            it = [0]

            def nxt():
                it[0] = (it[0] + 1) % copies
                return ws[it[0]]

            res = {"op": "smallm", "case": name, "B": B, "N": N, "K": K}
            ref = x.float() @ ws[0].float().t()
            got = gemm_smallm(x, ws[0])
            res["max_err"] = float((got.float() - ref).abs().max())
            res["ours_ms"] = graph_time_ms(lambda: gemm_smallm(x, nxt()))
            for sp in (1, 2, 4, 8):
                if sp * 2 <= (K // 64):
                    res[f"ours_s{sp}_ms"] = graph_time_ms(lambda: gemm_smallm(x, nxt(), splits=sp))
            res["cublas_ms"] = graph_time_ms(lambda: torch.matmul(x, nxt().t()))
            if B <= 16 and skinny_gemm(x, ws[0]) is not None:
                res["prev_ms"] = graph_time_ms(lambda: skinny_gemm(x, nxt()))
            else:
                res["prev_ms"] = graph_time_ms(lambda: gemm(x, nxt(), "nt"))
            bytes_ = N * K * 2 + B * (N + K) * 2
            res["roofline_ms"] = bytes_ / PEAKS["hbm_gbs"] / 1e6
            res["frac_of_measured_hbm"] = res["roofline_ms"] / res["ours_ms"]
            res["vs_cublas"] = res["cublas_ms"] / res["ours_ms"]
            rows.append(res)
            print(json.dumps(res), flush=True)
        del ws


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="gemm,attn,bw")
    ap.add_argument("--md", default="")
    args = ap.parse_args()
    assert _lib.available()
    rows = []
    if "gemm" in args.only:
        bench_gemm(rows)
    if "attn" in args.only:
        bench_attention(rows)
    if "bw" in args.only:
        bench_elementwise(rows)
    if "smallm" in args.only:
        bench_smallm(rows)
    if args.md:
        with open(args.md, "w") as f:
            f.write("| op | case | ours | library | ratio / roofline |\n|---|---|---|---|---|\n")
            for r in rows:
                if r["op"] == "gemm":
                    f.write(f"| gemm {r['layout']} | {r['case']} ({r['M']}x{r['N']}x{r['K']}) | {r['ours_ms']:.3f} ms, "
                            f"{r['ours_tflops']:.0f} TF | cuBLAS {r['cublas_ms']:.3f} ms, {r['cublas_tflops']:.0f} TF | "
                            f"{r['vs_cublas']:.2f}x cuBLAS, {r['frac_of_measured_peak']:.2f} of measured peak |\n")
                elif r["op"] == "smallm":
                    f.write(f"| gemm small-M | {r['case']} B={r['B']} ({r['N']}x{r['K']}) | {r['ours_ms'] * 1e3:.1f} us | cuBLAS "
                            f"{r['cublas_ms'] * 1e3:.1f} us, previous path {r['prev_ms'] * 1e3:.1f} us | {r['vs_cublas']:.2f}x "
                            f"cuBLAS, {r['frac_of_measured_hbm']:.2f} of measured HBM |\n")
                elif r["op"] == "flash_attn":
                    f.write(f"| flash_attn | {r['case']} | fwd {r['nxd_fwd_ms']:.3f} ms ({r['nxd_fwd_tflops']:.0f} TF), "
                            f"bwd {r['nxd_bwd_ms']:.3f} ms ({r['nxd_bwd_tflops']:.0f} TF) | SDPA fwd {r['sdpa_fwd_ms']:.3f}"
                            f" ms, bwd {r['sdpa_bwd_ms']:.3f} ms | fwd {r['fwd_frac_of_measured_peak']:.2f} of measured peak |\n")
                else:
                    f.write(f"| {r['case']} | | {r['ms']:.3f} ms, {r['gbs']:.0f} GB/s | | "
                            f"{r['frac_of_measured_hbm']:.2f} of measured HBM |\n")


if __name__ == "__main__":
    main()
