"""Python side of the fused compute+collective kernels.

``fused_ag_gemm``  all-gather(rows) -> GEMM in one kernel (``csrc/fused_ag_gemm.cu``): the kernel's copy clusters pull
the peers' shards over NVLink with the TMA engine while the tcgen05 clusters already multiply the slabs that arrived.
Returns ``None`` when the shape does not fit the kernel's constraints so that the caller takes the NCCL + GEMM path.
"""
from __future__ import annotations

import ctypes as _c
import os
from typing import Optional, Tuple

import torch

from . import _lib

_LAYOUT = {"nt": 0, "nn": 1}


def _comm_ctas() -> int:
    return int(os.environ.get("NXD_AG_COMM_CTAS", "12"))


def fused_ag_gemm(ctx, x_shard: torch.Tensor, weight: torch.Tensor, layout: str = "nt"
                  ) -> Optional[Tuple[torch.Tensor, torch.Tensor]]:
    """``(all_gather_rows(x_shard) @ W_op, x_total)`` or ``None`` if unsupported.  ``x_shard [R, K]`` bf16."""
    if layout not in _LAYOUT or x_shard.dtype != torch.bfloat16 or weight.dtype != torch.bfloat16:
        return None
    R, K = x_shard.shape
    N = weight.shape[0] if layout == "nt" else weight.shape[1]
    tp = ctx.world
    if R % 256 or K % 64 or N % 8 or tp > 8 or weight.stride(1) != 1 or weight.stride(0) % 8:
        return None
    x_shard = x_shard.contiguous()
    shard_bytes = R * K * 2
    # symmetric shard slots: two per rank (double buffering, see the kernel's header comment)
    _, table = ctx.region("ag_slots", 2 * shard_bytes)
    chunks = R // 256
    if not hasattr(ctx, "_ag_state"):
        ctx._ag_state = {"ready": ctx.signal_slots("ag_ready", 8), "own": ctx.signal_slots("ag_own", 1),
                         "chunks": ctx.signal_slots("ag_chunks", 8 * 64), "host_tables": {}}
    st = ctx._ag_state
    assert chunks <= 64, "rows per rank too large for the chunk flag table"
    epoch = ctx.next_epoch("ag_gemm")
    parity = epoch & 1
    key = (table.data_ptr(), shard_bytes, parity)
    if key not in st["host_tables"]:
        peers = table.tolist()
        slots = (_c.c_void_p * tp)(*[p + parity * shard_bytes for p in peers])
        sigs = (_c.c_void_p * tp)(*ctx.sig_peers)
        st["host_tables"][key] = (slots, sigs)
    slots, sigs = st["host_tables"][key]
    # NOTE: the chunk counters are indexed [src * chunks_per_slab + c]; they count copy CTAs across calls, so the
    # number of copy CTAs and the chunk count must not change between calls on one context.
    cfg = (chunks, _comm_ctas())
    if st.setdefault("cfg", cfg) != cfg:
        return None
    x_total = torch.empty((tp * R, K), dtype=torch.bfloat16, device=x_shard.device)
    y = torch.empty((tp * R, N), dtype=torch.bfloat16, device=x_shard.device)
    rc = _lib.lib().nxd_ag_gemm_bf16(
        _c.c_int(_LAYOUT[layout]), _lib.ptr(x_shard), _lib.ptr(x_total), _lib.ptr(weight), _lib.ptr(y), _c.c_int(R),
        _c.c_int(N), _c.c_int(K), _c.c_longlong(weight.stride(0)), _c.c_longlong(y.stride(0)), slots, sigs,
        _c.c_int(tp), _c.c_int(ctx.rank), _c.c_int(st["ready"]), _c.c_int(st["chunks"]), _c.c_int(st["own"]),
        _c.c_uint(epoch), _c.c_int(cfg[1]), _c.c_int(0), _lib.stream_ptr())
    _lib.check(rc, "nxd_ag_gemm_bf16")
    return y, x_total


def fused_gemm_rs(ctx, x2d, weight, layout="nt"):
    return None


def fused_gemm_ar(ctx, x2d, weight, layout="nt"):
    return None
