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


def _comm_ctas(tp: int) -> int:
    """Copy-engine CTAs of the AG->GEMM kernel (even).  More ranks = more NVLink traffic per unit of math."""
    env = os.environ.get("NXD_AG_COMM_CTAS")
    if env:
        return int(env)
    return {2: 12, 4: 16}.get(tp, 24)


_trace_on = os.environ.get("NXD_FUSED_TRACE", "0") == "1"
_traces = []


def _trace_buf(device, what: str):
    if not _trace_on:
        return None
    t = torch.zeros(16, dtype=torch.int64, device=device)
    _traces.append((what, t))
    return t


def dump_traces(limit: int = 6):  # noqa: D401
    """Debug: print the globaltimer stamps (us, relative to the first) collected with NXD_FUSED_TRACE=1."""
    torch.cuda.synchronize()
    out = []
    for what, t in _traces[-limit:]:
        v = t.tolist()
        base = min(x for x in v if x > 0) if any(x > 0 for x in v) else 0
        out.append((what, [round((x - base) / 1e3, 1) if x > 0 else None for x in v[:12]]))
    return out


def _ag_block_n(tp: int, R: int, N: int, comm_ctas: int) -> int:
    """Tile width of the AG->GEMM kernel.  256 everywhere: at TP8 the 128-wide tiles were tried for the narrow
    projections (a wave-count model picked them for qkv N = 640 and gate_up N = 2048) and lost -- 123 vs 115 us and
    171 vs 131 us (``profiles/r2/fused_comm_tp8_bn*.jsonl``): the kernel is bound by the gather, and narrower tiles only
    add B-operand traffic.  ``NXD_AG_BLOCK_N=128`` keeps the variant reachable (the tests run both)."""
    env = os.environ.get("NXD_AG_BLOCK_N")
    return int(env) if env else 256


def _host_ptrs(ctx, table: torch.Tensor, offset: int):
    return (_c.c_void_p * ctx.world)(*[p + offset for p in table.tolist()])


def fused_ag_gemm(ctx, x_shard: torch.Tensor, weight: torch.Tensor, layout: str = "nt", persistent_total: bool = True
                  ) -> Optional[Tuple[torch.Tensor, torch.Tensor]]:
    """``(all_gather_rows(x_shard) @ W_op, x_total)`` or ``None`` if unsupported.  ``x_shard [R, K]`` bf16.
    ``x_total`` lives in a double-buffered symmetric region: with ``persistent_total`` it is copied to a private
    tensor (needed when it is kept for backward), otherwise the caller must consume it before the second next call."""
    if layout not in _LAYOUT or x_shard.dtype != torch.bfloat16 or weight.dtype != torch.bfloat16:
        return None
    R, K = x_shard.shape
    N = weight.shape[0] if layout == "nt" else weight.shape[1]
    tp = ctx.world
    if R % 256 or K % 64 or N % 8 or tp > 8 or weight.stride(1) != 1 or weight.stride(0) % 8 or R // 256 > 64:
        return None
    x_shard = x_shard.contiguous()
    total_bytes = tp * R * K * 2
    comm_ctas = _comm_ctas(tp)
    # counters / buffers are per configuration: the counters accumulate a fixed amount per call
    cfg = f"ag:{R}:{K}:{comm_ctas}"
    mc_base = 0
    # NVLS: the copy CTAs store the local shard ONCE to the multicast address of the gathered buffer (multimem.st, the
    # switch replicates it) instead of once per peer.  Verified at 2 and 8 GPUs; TP8 step 55.8 -> 53.2 ms
    # (profiles/r2/bench_tp8_nvls.json).  Falls back to per-peer pushes where no multicast object can be bound.
    nvls = os.environ.get("NXD_NVLS")          # unset: on for tp > 2 (one peer gains nothing); "1": always; "0": never
    if nvls == "1" or (nvls is None and tp > 2):
        got = ctx.region_mc(cfg, 2 * total_bytes)
        if got is not None:
            local_ptr, table, mc_base = got
    if not mc_base:
        local_ptr, table = ctx.region(cfg, 2 * total_bytes)   # X_total, double buffered
    st = ctx.__dict__.setdefault("_fused_state", {})
    if cfg not in st:
        st[cfg] = {"chunks": ctx.signal_slots(cfg + ":chunks", 8 * (R // 256)),
                   "sigs": (_c.c_void_p * tp)(*ctx.sig_peers), "bufs": {}}
    s = st[cfg]
    epoch = ctx.next_epoch(cfg)
    parity = epoch & 1
    key = (table.data_ptr(), parity)
    if key not in s["bufs"]:
        s["bufs"][key] = (_host_ptrs(ctx, table, parity * total_bytes),
                          ctx.tensor_view(local_ptr + parity * total_bytes, (tp * R, K), torch.bfloat16))
    xt_ptrs, x_total = s["bufs"][key]
    y = torch.empty((tp * R, N), dtype=torch.bfloat16, device=x_shard.device)
    if mc_base:
        rc = _lib.lib().nxd_ag_gemm_bf16_nvls(
            _c.c_int(_LAYOUT[layout]), _lib.ptr(x_shard), _lib.ptr(weight), _lib.ptr(y), _c.c_int(R), _c.c_int(N),
            _c.c_int(K), _c.c_longlong(weight.stride(0)), _c.c_longlong(y.stride(0)), xt_ptrs, s["sigs"], _c.c_int(tp),
            _c.c_int(ctx.rank), _c.c_int(s["chunks"]), _c.c_uint(epoch), _c.c_int(comm_ctas), _c.c_int(0),
            _c.c_void_p(mc_base + parity * total_bytes), _lib.stream_ptr())
        _lib.check(rc, "nxd_ag_gemm_bf16_nvls")
        return y, (x_total.clone() if persistent_total else x_total)
    trace = _trace_buf(x_shard.device, f"ag {R}x{K}x{N}")
    rc = _lib.lib().nxd_ag_gemm_bf16(
        _c.c_int(_LAYOUT[layout]), _lib.ptr(x_shard), _lib.ptr(weight), _lib.ptr(y), _c.c_int(R), _c.c_int(N),
        _c.c_int(K), _c.c_longlong(weight.stride(0)), _c.c_longlong(y.stride(0)), xt_ptrs, s["sigs"], _c.c_int(tp),
        _c.c_int(ctx.rank), _c.c_int(s["chunks"]), _c.c_uint(epoch), _c.c_int(comm_ctas),
        _c.c_int(_ag_block_n(tp, R, N, comm_ctas)), _lib.ptr(trace), _lib.stream_ptr())
    _lib.check(rc, "nxd_ag_gemm_bf16")
    return y, (x_total.clone() if persistent_total else x_total)


def fused_gemm_rs(ctx, x2d: torch.Tensor, weight: torch.Tensor, layout: str = "nt") -> Optional[torch.Tensor]:
    """``reduce_scatter_rows(x2d @ W_op)`` in one kernel (``csrc/fused_gemm_rs.cu``) or ``None`` if unsupported.
    ``x2d [tp*R, Kloc]`` bf16; returns this rank's ``[R, N]`` rows of the sum over ranks."""
    if layout not in _LAYOUT or x2d.dtype != torch.bfloat16 or weight.dtype != torch.bfloat16:
        return None
    tp = ctx.world
    M, K = x2d.shape
    N = weight.shape[0] if layout == "nt" else weight.shape[1]
    if M % tp:
        return None
    R = M // tp
    if (R % 256 or N % 8 or K % 8 or tp > 8 or weight.stride(1) != 1 or weight.stride(0) % 8 or x2d.stride(1) != 1
            or x2d.stride(0) % 8 or R // 256 > 64):
        return None
    bn = 256
    cfg = f"rs:{R}:{N}:{bn}"
    stage_bytes = tp * R * N * 2
    _, table = ctx.region(cfg, 2 * stage_bytes)
    st = ctx.__dict__.setdefault("_fused_state", {})
    if cfg not in st:
        st[cfg] = {"cnt": ctx.signal_slots(cfg + ":cnt", 8 * (R // 256)), "sigs": (_c.c_void_p * tp)(*ctx.sig_peers),
                   "stages": {}}
    s = st[cfg]
    epoch = ctx.next_epoch(cfg)
    parity = epoch & 1
    key = (table.data_ptr(), parity)
    if key not in s["stages"]:
        s["stages"][key] = _host_ptrs(ctx, table, parity * stage_bytes)
    out = torch.empty((R, N), dtype=torch.bfloat16, device=x2d.device)
    rc = _lib.lib().nxd_gemm_rs_bf16(
        _c.c_int(_LAYOUT[layout]), _lib.ptr(x2d), _lib.ptr(weight), _lib.ptr(out), _c.c_int(R), _c.c_int(N), _c.c_int(K),
        _c.c_longlong(x2d.stride(0)), _c.c_longlong(weight.stride(0)), _c.c_longlong(out.stride(0)), s["stages"][key],
        s["sigs"], _c.c_int(tp), _c.c_int(ctx.rank), _c.c_int(s["cnt"]), _c.c_uint(epoch), _c.c_int(bn),
        _lib.ptr(_trace_buf(x2d.device, f"rs {R}x{K}x{N}")), _lib.stream_ptr())
    _lib.check(rc, "nxd_gemm_rs_bf16")
    return out


# ---- decode path: one-shot all-reduce and skinny GEMM -> all-reduce (csrc/fused_gemm_ar.cu) -----------------------------
_AR_SLOT_ELEMS = 1 << 20           # floats per (parity, source) staging slot: messages up to 2 MB of bf16
_SKINNY_MAX_ROWS = 16             # 1..8 rows: FMA kernel; 9..16: mma.sync m16n8k16 kernel (same entry point)


def _ar_state(ctx):
    st = ctx.__dict__.setdefault("_fused_state", {})
    if "ar" not in st:
        tp = ctx.world
        _, table = ctx.region("ar", 2 * tp * _AR_SLOT_ELEMS * 4)
        st["ar"] = {"slot": ctx.signal_slots("ar:flags", int(_lib.lib().nxd_oneshot_num_flags())),
                    "sigs": (_c.c_void_p * tp)(*ctx.sig_peers), "stages": _host_ptrs(ctx, table, 0),
                    # call counter + CTA counter on the device: launches stay valid when replayed from a CUDA graph
                    "state": torch.zeros(2, dtype=torch.int32, device=ctx.device)}
    return st["ar"]


def oneshot_all_reduce(ctx, x: torch.Tensor, out: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
    """Sum of ``x`` (bf16, contiguous, numel % 8 == 0, <= 1 Mi elements) over the ranks of ``ctx`` in ONE kernel: every
    rank pushes its values into all peers' staging slots and adds the tp slots in rank order (fp32).  ``None`` when
    the message does not qualify (caller falls back to NCCL)."""
    n = x.numel()
    if x.dtype != torch.bfloat16 or not x.is_contiguous() or n % 8 or n == 0 or ctx.world > 8:
        return None
    per_cta = 2048
    while (n + per_cta - 1) // per_cta > 132:
        per_cta *= 2
    if per_cta * 4 > 160 * 1024 or ((n + per_cta - 1) // per_cta) * per_cta > _AR_SLOT_ELEMS:
        return None
    s = _ar_state(ctx)
    out = torch.empty_like(x) if out is None else out
    rc = _lib.lib().nxd_allreduce_oneshot(_lib.ptr(x), _lib.ptr(out), _c.c_longlong(n), s["stages"], s["sigs"],
                                          _c.c_size_t(_AR_SLOT_ELEMS), _c.c_int(ctx.world), _c.c_int(ctx.rank),
                                          _c.c_int(s["slot"]), _lib.ptr(s["state"]), _lib.stream_ptr())
    _lib.check(rc, "nxd_allreduce_oneshot")
    return out


def skinny_gemm(x2d: torch.Tensor, weight: torch.Tensor, residual: Optional[torch.Tensor] = None, ctx=None
                ) -> Optional[torch.Tensor]:
    """``x2d[B, K] @ weight[N, K]^T (+ residual)`` for B <= 16 rows, summed over the ranks of ``ctx`` when given: the
    weight-streaming kernel of the token-generation path.  ``None`` when the shape does not qualify."""
    B, K = x2d.shape
    N = weight.shape[0]
    if (B > _SKINNY_MAX_ROWS or B == 0 or x2d.dtype != torch.bfloat16 or weight.dtype != torch.bfloat16 or K % 8 or N % 4
            or x2d.stride(1) != 1 or weight.stride(1) != 1 or x2d.stride(0) % 8 or weight.stride(0) % 8):
        return None
    kb = 1 if B <= 1 else 2 if B <= 2 else 4 if B <= 4 else 8 if B <= 8 else 16
    if kb * (K + 8) * 2 + kb * 64 * 4 + 4096 > 160 * 1024:
        return None
    tp = 1 if ctx is None else ctx.world
    if tp > 8 or (tp > 1 and kb * (N + 8192) > _AR_SLOT_ELEMS):
        return None
    if tp > 1 and B > 8:
        return None     # the 16-row kernel's all-reduce epilogue has only been run single-GPU: GEMM + one-shot AR instead
    out = torch.empty((B, N), dtype=torch.bfloat16, device=x2d.device)
    if residual is not None:
        residual = residual.reshape(B, N)
        if residual.stride(1) != 1 or residual.stride(0) != out.stride(0):
            residual = residual.contiguous()
    if tp > 1:
        s = _ar_state(ctx)
        stages, sigs, slot, rank, state = s["stages"], s["sigs"], s["slot"], ctx.rank, s["state"]
    else:
        stages = sigs = state = None
        slot = rank = 0
    rc = _lib.lib().nxd_skinny_gemm_ar_bf16(
        _lib.ptr(x2d), _lib.ptr(weight), _lib.ptr(residual), _lib.ptr(out), _c.c_int(B), _c.c_int(N), _c.c_int(K),
        _c.c_longlong(x2d.stride(0)), _c.c_longlong(weight.stride(0)), _c.c_longlong(out.stride(0)), stages, sigs,
        _c.c_size_t(_AR_SLOT_ELEMS), _c.c_int(tp), _c.c_int(rank), _c.c_int(slot), _lib.ptr(state), _lib.stream_ptr())
    _lib.check(rc, "nxd_skinny_gemm_ar_bf16")
    return out


def skinny_gemm_i8(x2d: torch.Tensor, weight_i8: torch.Tensor, scale: torch.Tensor,
                   residual: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
    """``(x2d[B, K] @ weight_i8[N, K]^T) * scale (+ residual)`` for B <= 8 bf16 rows and int8 weights: the weight-only
    quantized token-generation GEMM (``csrc/skinny_int8.cu``).  ``scale`` has 1 element (per tensor) or N (per output
    channel).  ``None`` when the shape does not qualify."""
    B, K = x2d.shape
    N = weight_i8.shape[0]
    sn = scale.numel()
    if (B == 0 or B > 8 or x2d.dtype != torch.bfloat16 or weight_i8.dtype != torch.int8 or K % 16 or x2d.stride(1) != 1
            or weight_i8.stride(1) != 1 or x2d.stride(0) % 8 or weight_i8.stride(0) % 16 or sn not in (1, N)):
        return None
    kb = 1 if B <= 1 else 2 if B <= 2 else 4 if B <= 4 else 8
    if kb * K * 2 + kb * 16 * 4 > 200 * 1024:
        return None
    sc = scale.reshape(-1).to(torch.float32).contiguous()
    out = torch.empty((B, N), dtype=torch.bfloat16, device=x2d.device)
    if residual is not None:
        residual = residual.reshape(B, N)
        if residual.stride(1) != 1 or residual.stride(0) != out.stride(0):
            residual = residual.contiguous()
    rc = _lib.lib().nxd_skinny_gemm_i8(
        _lib.ptr(x2d), _lib.ptr(weight_i8), _lib.ptr(sc), _c.c_int(sn), _lib.ptr(residual), _lib.ptr(out), _c.c_int(B),
        _c.c_int(N), _c.c_int(K), _c.c_longlong(x2d.stride(0)), _c.c_longlong(weight_i8.stride(0)),
        _c.c_longlong(out.stride(0)), _lib.stream_ptr())
    _lib.check(rc, "nxd_skinny_gemm_i8")
    return out


def fused_gemm_ar(ctx, x2d: torch.Tensor, weight: torch.Tensor, layout: str = "nt") -> Optional[torch.Tensor]:
    """``all_reduce(x2d @ W_op)``: one kernel for token generation (<= 16 rows), tcgen05 GEMM + one-shot all-reduce
    for messages up to 2 MB (short prefill), ``None`` beyond (NCCL's ring / NVLS wins on bandwidth there)."""
    if x2d.dtype != torch.bfloat16 or weight.dtype != torch.bfloat16:
        return None
    if layout == "nt" and x2d.shape[0] <= _SKINNY_MAX_ROWS:
        res = skinny_gemm(x2d, weight, ctx=ctx)
        if res is not None:
            return res
    N = weight.shape[0] if layout == "nt" else weight.shape[1]
    if x2d.shape[0] * N > _AR_SLOT_ELEMS or (x2d.shape[0] * N) % 8:
        return None
    from . import linear as gemm_ops

    y = gemm_ops.gemm_nt(x2d, weight) if layout == "nt" else gemm_ops.gemm_nn(x2d, weight)
    return oneshot_all_reduce(ctx, y.contiguous(), out=y if y.is_contiguous() else None)


# ---- expert parallelism: dispatch / combine all-to-all fused with the expert GEMMs (csrc/fused_ag_gemm.cu, fused_gemm_rs.cu) ---
def _expert_weight_2d(weight: torch.Tensor, layout: str) -> Optional[torch.Tensor]:
    """Stacked expert weights ``[E_loc, a, b]`` as the 2-D operand of the fused kernels (``nn``: each expert is
    ``[K, N]``; ``nt``: each expert is ``[N, K]``)."""
    if weight.dim() != 3 or not weight.is_contiguous() or weight.dtype != torch.bfloat16:
        return None
    return weight.reshape(weight.shape[0] * weight.shape[1], weight.shape[2])


def fused_a2a_gemm(ctx, x: torch.Tensor, weight: torch.Tensor, layout: str = "nn"
                   ) -> Optional[Tuple[torch.Tensor, torch.Tensor]]:
    """Expert-parallel dispatch + expert GEMM in one kernel.  ``x [ep, E_loc, C, K]``: slab ``d`` = tokens routed to
    the experts of EP rank ``d``.  Returns ``(y [ep(src), E_loc, C, N], x_total [ep(src), E_loc, C, K])`` -- the
    received tokens are a by-product kept for the weight gradient (a private copy) -- or ``None`` if unsupported."""
    if layout not in _LAYOUT or x.dtype != torch.bfloat16 or x.dim() != 4:
        return None
    ep, e_loc, C, K = x.shape
    w2d = _expert_weight_2d(weight, layout)
    if w2d is None or ep != ctx.world or ep > 8 or weight.shape[0] != e_loc:
        return None
    N = weight.shape[2] if layout == "nn" else weight.shape[1]
    if (weight.shape[1] if layout == "nn" else weight.shape[2]) != K:
        return None
    R = e_loc * C
    if C % 256 or K % 64 or N % 8 or R // 256 > 64:
        return None
    x = x.contiguous()
    total_bytes = ep * R * K * 2
    comm_ctas = _comm_ctas(ep)
    cfg = f"a2a:{R}:{K}:{comm_ctas}"
    local_ptr, table = ctx.region(cfg, 2 * total_bytes)
    st = ctx.__dict__.setdefault("_fused_state", {})
    if cfg not in st:
        st[cfg] = {"chunks": ctx.signal_slots(cfg + ":chunks", 8 * (R // 256)),
                   "sigs": (_c.c_void_p * ep)(*ctx.sig_peers), "bufs": {}}
    s = st[cfg]
    epoch = ctx.next_epoch(cfg)
    parity = epoch & 1
    key = (table.data_ptr(), parity)
    if key not in s["bufs"]:
        s["bufs"][key] = (_host_ptrs(ctx, table, parity * total_bytes),
                          ctx.tensor_view(local_ptr + parity * total_bytes, (ep, e_loc, C, K), torch.bfloat16))
    xt_ptrs, x_total = s["bufs"][key]
    y = torch.empty((ep, e_loc, C, N), dtype=torch.bfloat16, device=x.device)
    rc = _lib.lib().nxd_a2a_gemm_bf16(
        _c.c_int(_LAYOUT[layout]), _lib.ptr(x), _lib.ptr(w2d), _lib.ptr(y), _c.c_int(R), _c.c_int(C), _c.c_int(N),
        _c.c_int(K), _c.c_longlong(w2d.stride(0)), _c.c_longlong(N), xt_ptrs, s["sigs"], _c.c_int(ep),
        _c.c_int(ctx.rank), _c.c_int(s["chunks"]), _c.c_uint(epoch), _c.c_int(comm_ctas), _c.c_int(0), None,
        _lib.stream_ptr())
    _lib.check(rc, "nxd_a2a_gemm_bf16")
    return y, x_total.clone()


def fused_gemm_a2a(ctx, x: torch.Tensor, weight: torch.Tensor, layout: str = "nn") -> Optional[torch.Tensor]:
    """Expert GEMM + expert-parallel combine in one kernel.  ``x [ep(src), E_loc, C, K]`` (slab ``s`` = tokens that came
    from EP rank ``s``); every result tile is stored straight into the output of the rank the tokens came from.
    Returns this rank's ``[E, C, N]`` (``E = ep * E_loc``, expert-major as before the dispatch) or ``None``."""
    if layout not in _LAYOUT or x.dtype != torch.bfloat16 or x.dim() != 4:
        return None
    ep, e_loc, C, K = x.shape
    w2d = _expert_weight_2d(weight, layout)
    if w2d is None or ep != ctx.world or ep > 8 or weight.shape[0] != e_loc:
        return None
    N = weight.shape[2] if layout == "nn" else weight.shape[1]
    if (weight.shape[1] if layout == "nn" else weight.shape[2]) != K:
        return None
    R = e_loc * C
    if C % 256 or K % 8 or N % 8 or R // 256 > 64:
        return None
    x = x.contiguous()
    bn = 256
    cfg = f"ga2a:{R}:{N}:{bn}"
    out_bytes = ep * R * N * 2
    local_ptr, table = ctx.region(cfg, 2 * out_bytes)
    st = ctx.__dict__.setdefault("_fused_state", {})
    if cfg not in st:
        st[cfg] = {"cnt": ctx.signal_slots(cfg + ":cnt", 8 * (R // 256)), "sigs": (_c.c_void_p * ep)(*ctx.sig_peers),
                   "outs": {}}
    s = st[cfg]
    epoch = ctx.next_epoch(cfg)
    parity = epoch & 1
    key = (table.data_ptr(), parity)
    if key not in s["outs"]:
        s["outs"][key] = (_host_ptrs(ctx, table, parity * out_bytes),
                          ctx.tensor_view(local_ptr + parity * out_bytes, (ep * e_loc, C, N), torch.bfloat16))
    out_ptrs, out = s["outs"][key]
    x2d = x.reshape(ep * R, K)
    rc = _lib.lib().nxd_gemm_a2a_bf16(
        _c.c_int(_LAYOUT[layout]), _lib.ptr(x2d), _lib.ptr(w2d), _c.c_int(R), _c.c_int(C), _c.c_int(N), _c.c_int(K),
        _c.c_longlong(x2d.stride(0)), _c.c_longlong(w2d.stride(0)), out_ptrs, s["sigs"], _c.c_int(ep),
        _c.c_int(ctx.rank), _c.c_int(s["cnt"]), _c.c_uint(epoch), _c.c_int(bn), None, _lib.stream_ptr())
    _lib.check(rc, "nxd_gemm_a2a_bf16")
    return out.clone()      # the symmetric buffer is rewritten two calls later
