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
torch.cuda.synchronize()
print("done")
