#!/usr/bin/env python
"""Tiny driver for `ncu --set full`: runs each hot kernel a few times at the Llama-3.2-3B (tp1, S=8192) shapes."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from neuronx_distributed_llama3_2_b200.ops import attention as A  # noqa: E402
from neuronx_distributed_llama3_2_b200.ops.gemm import gemm  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "all"
torch.manual_seed(0)
if which in ("all", "gemm"):
    a = torch.randn(8192, 3072, device="cuda").bfloat16()
    w = torch.randn(16384, 3072, device="cuda").bfloat16()
    for _ in range(3):
        gemm(a, w, "nt")
if which in ("all", "attn"):
    q = torch.randn(1, 24, 8192, 128, device="cuda").bfloat16().requires_grad_()
    k = torch.randn(1, 8, 8192, 128, device="cuda").bfloat16().requires_grad_()
    v = torch.randn(1, 8, 8192, 128, device="cuda").bfloat16().requires_grad_()
    do = torch.randn(1, 24, 8192, 128, device="cuda").bfloat16()
    for _ in range(2):
        o = A.flash_attn_func(q, k, v, causal=True)
        o.backward(do)
if which in ("all", "bw"):
    from neuronx_distributed_llama3_2_b200.ops import activation, cross_entropy as ce, norm, optim

    T, H, I = 8192, 3072, 8192
    x = torch.randn(T, H, device="cuda").bfloat16().requires_grad_()
    r = torch.randn(T, H, device="cuda").bfloat16()
    w = torch.ones(H, device="cuda").bfloat16().requires_grad_()
    y = norm.rms_norm(x, w, 1e-5)
    y.backward(torch.randn_like(y))
    gu = torch.randn(T, 2 * I, device="cuda").bfloat16().requires_grad_()
    a = activation.swiglu(gu)
    a.backward(torch.randn_like(a))
    n = 100_000_000
    master = torch.randn(n, device="cuda"); m = torch.zeros(n, device="cuda"); v2 = torch.zeros(n, device="cuda")
    g = torch.randn(n, device="cuda").bfloat16(); po = torch.empty(n, device="cuda", dtype=torch.bfloat16)
    optim.adamw_step_(master, m, v2, g, po, 1e-3, 0.9, 0.95, 1e-8, 0.1, 1)
    V = 128256
    logits = torch.randn(2048, V, device="cuda").bfloat16()
    tgt = torch.randint(0, V, (2048,), device="cuda")
    ce.ce_local_stats(logits, tgt, 0)
if which in ("all", "decode"):
    from neuronx_distributed_llama3_2_b200.ops import decode_fused, rope as rope_ops
    from neuronx_distributed_llama3_2_b200.ops.fused_comm import skinny_gemm

    B, H, nq, nkv, D, S, I = 1, 2048, 32, 8, 64, 4096, 8192
    xd = torch.randn(B, H, device="cuda").bfloat16()
    nw = torch.ones(H, device="cuda").bfloat16()
    wqkv = (torch.randn((nq + 2 * nkv) * D, H, device="cuda") * 0.02).bfloat16()
    cos, sin = rope_ops.build_rope_cache(S, D, 500000.0, None, device="cuda")
    ck = torch.randn(16, nkv, S, D, device="cuda").bfloat16()
    cv = torch.randn(16, nkv, S, D, device="cuda").bfloat16()
    pos = torch.full((B,), 2048, device="cuda", dtype=torch.int32)
    for _ in range(2):
        qd = decode_fused.qkv_rope_cache(xd, nw, 1e-5, wqkv, cos, sin, pos, ck, cv, nq, nkv, D)
        o = A.decode_attention(qd, ck, cv, pos + 1)
        wgu = (torch.randn(2 * I, H, device="cuda") * 0.02).bfloat16()
        act = decode_fused.gate_up_swiglu(xd, nw, 1e-5, wgu)
        wd = (torch.randn(H, I, device="cuda") * 0.02).bfloat16()
        skinny_gemm(act, wd, residual=xd)
    q16 = torch.randn(16, nq, D, device="cuda").bfloat16()
    A.decode_attention(q16, ck, cv, torch.full((16,), 2049, device="cuda", dtype=torch.int32))
if which in ("all", "decode_batched"):
    # batched token generation (Llama-3.2-1B, 64 sequences at position 2048): the small-M tensor-core GEMM family and
    # the tensor-core decode attention.  Weights are cycled so that every launch streams them from HBM.
    from neuronx_distributed_llama3_2_b200.ops import decode_fused, rope as rope_ops
    from neuronx_distributed_llama3_2_b200.ops.gemm import gemm_smallm

    B, H, nq, nkv, D, S, I = 64, 2048, 32, 8, 64, 4096, 8192
    xn = torch.randn(B, H, device="cuda").bfloat16()
    cos, sin = rope_ops.build_rope_cache(S, D, 500000.0, None, device="cuda")
    ck = torch.randn(B, nkv, S, D, device="cuda").bfloat16()
    cv = torch.randn(B, nkv, S, D, device="cuda").bfloat16()
    pos = torch.full((B,), 2048, device="cuda", dtype=torch.int32)
    wq = [(torch.randn((nq + 2 * nkv) * D, H, device="cuda") * 0.02).bfloat16() for _ in range(24)]
    wo = [(torch.randn(H, nq * D, device="cuda") * 0.02).bfloat16() for _ in range(36)]
    wg = [(torch.randn(2 * I, H, device="cuda") * 0.02).bfloat16() for _ in range(6)]
    wd = [(torch.randn(H, I, device="cuda") * 0.02).bfloat16() for _ in range(10)]
    for i in range(max(len(wq), len(wo))):      # last launches of each kernel see cold weights
        qd = decode_fused.qkv_rope_cache_batched(xn, wq[i % len(wq)], cos, sin, pos, ck, cv, nq, nkv, D)
        o = A.decode_attention(qd, ck, cv, pos + 1)
        x2 = gemm_smallm(o.reshape(B, nq * D), wo[i % len(wo)], residual=xn)
        act = decode_fused.gate_up_swiglu_batched(x2, wg[i % len(wg)])
        gemm_smallm(act, wd[i % len(wd)], residual=x2)
torch.cuda.synchronize()
print("done")
