#!/usr/bin/env python
"""Phase timeline of CTA 0 of the small-M GEMM (library built with -DNXD_SMALLM_TRACE): SM-clock stamps -> ns deltas."""
import ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from neuronx_distributed_llama3_2_b200.ops import _lib
from neuronx_distributed_llama3_2_b200.ops.gemm import gemm_smallm

names = ["entry", "setup done", "pdl wait done", "first tile landed", "last mma issued", "acc ready (epi)", "pushed",
         "cluster sync", "owner pass done", "exit"]
mhz = 1900.0
for (B, N, K, sp) in [(16, 2048, 2048, 8), (64, 2048, 2048, 8), (64, 2048, 2048, 1), (64, 16384, 2048, 1), (64, 2048, 8192, 8)]:
    x = torch.randn(B, K, device="cuda").bfloat16()
    ws = [torch.randn(N, K, device="cuda").bfloat16() for _ in range(max(2, (300 << 20) // (N * K * 2)))]
    for i in range(5):
        gemm_smallm(x, ws[i % len(ws)], splits=sp)
    torch.cuda.synchronize()
    # isolated launch (nothing before it in the stream)
    gemm_smallm(x, ws[-1], splits=sp)
    torch.cuda.synchronize()
    buf = (ctypes.c_longlong * 16)()
    _lib.lib().nxd_smallm_trace(buf)
    t = list(buf)[:10]
    print(f"B={B} N={N} K={K} splits={sp}: " + " | ".join(f"{names[i]} +{(t[i] - t[0]) / mhz * 1e3:.0f}ns" for i in range(1, 10)))
