"""Vocab-parallel embedding gather."""
from __future__ import annotations

import ctypes as _c

import torch

from . import _lib


class _VocabParallelEmbeddingFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ids, weight, vocab_start, seq_first_out):
        B, S = ids.shape if ids.dim() == 2 else (1, ids.numel())
        H = weight.shape[1]
        ids_c = ids.reshape(-1).to(torch.int64).contiguous()
        T = ids_c.numel()
        out = torch.empty((T, H), dtype=weight.dtype, device=weight.device)
        rc = _lib.lib().nxd_embedding_fwd(_lib.ptr(ids_c), _lib.ptr(weight), _lib.ptr(out), _c.c_longlong(T),
                                          _c.c_int(H), _c.c_int(vocab_start), _c.c_int(weight.shape[0]), _c.c_int(B),
                                          _c.c_int(S), _c.c_int(1 if seq_first_out else 0), _lib.stream_ptr())
        _lib.check(rc, "nxd_embedding_fwd")
        ctx.save_for_backward(ids_c)
        ctx.meta = (vocab_start, weight.shape, weight.dtype, B, S, seq_first_out)
        return out.reshape((S, B, H) if seq_first_out else (B, S, H))

    @staticmethod
    def backward(ctx, dy):
        (ids_c,) = ctx.saved_tensors
        vocab_start, wshape, wdtype, B, S, seq_first = ctx.meta
        dy2 = dy.reshape(-1, wshape[1]).contiguous()
        dw = torch.zeros(wshape, dtype=torch.float32, device=dy.device)
        rc = _lib.lib().nxd_embedding_bwd(_lib.ptr(ids_c), _lib.ptr(dy2), _lib.ptr(dw), _c.c_longlong(ids_c.numel()),
                                          _c.c_int(wshape[1]), _c.c_int(vocab_start), _c.c_int(wshape[0]),
                                          _c.c_int(B), _c.c_int(S), _c.c_int(1 if seq_first else 0),
                                          _lib.stream_ptr())
        _lib.check(rc, "nxd_embedding_bwd")
        return None, dw.to(wdtype), None, None


class _SPEmbeddingPushFn(torch.autograd.Function):
    """Sequence-parallel embedding forward as one kernel over NVLink peer memory (``nxd_embedding_fwd_sp_push``):
    returns this rank's ``[S/tp, B, H]`` shard.  Backward = all-gather of dY over the TP group + the scatter-add
    kernel (what ``reduce_scatter_to_sequence_parallel_region`` + the embedding backward do on the NCCL path)."""

    @staticmethod
    def forward(ctx, ids, weight, vocab_start, sym):
        B, S = ids.shape
        H = weight.shape[1]
        tp = sym.world
        ids_c = ids.reshape(-1).to(torch.int64).contiguous()
        shard_bytes = (S // tp) * B * H * 2
        cfg = f"emb:{S}:{B}:{H}"
        local_ptr, table = sym.region(cfg, 2 * shard_bytes)
        st = sym.__dict__.setdefault("_fused_state", {})
        if cfg not in st:
            st[cfg] = {"slot": sym.signal_slots(cfg + ":flags", 8), "sigs": (_c.c_void_p * tp)(*sym.sig_peers),
                       "done": torch.zeros(1, dtype=torch.int32, device=weight.device), "outs": {}}
        s = st[cfg]
        epoch = sym.next_epoch(cfg)
        parity = epoch & 1
        key = (table.data_ptr(), parity)
        if key not in s["outs"]:
            s["outs"][key] = ((_c.c_void_p * tp)(*[p + parity * shard_bytes for p in table.tolist()]),
                              sym.tensor_view(local_ptr + parity * shard_bytes, (S // tp, B, H), torch.bfloat16))
        out_ptrs, view = s["outs"][key]
        rc = _lib.lib().nxd_embedding_fwd_sp_push(
            _lib.ptr(ids_c), _lib.ptr(weight), out_ptrs, s["sigs"], _lib.ptr(s["done"]), _c.c_longlong(ids_c.numel()),
            _c.c_int(H), _c.c_int(vocab_start), _c.c_int(weight.shape[0]), _c.c_int(B), _c.c_int(S), _c.c_int(tp),
            _c.c_int(sym.rank), _c.c_int(s["slot"]), _c.c_uint(epoch), _lib.stream_ptr())
        _lib.check(rc, "nxd_embedding_fwd_sp_push")
        ctx.save_for_backward(ids_c)
        ctx.meta = (vocab_start, weight.shape, weight.dtype, B, S)
        return view.clone()     # the symmetric buffer is reused two calls later

    @staticmethod
    def backward(ctx, dy):
        from ..parallel import mappings

        (ids_c,) = ctx.saved_tensors
        vocab_start, wshape, wdtype, B, S = ctx.meta
        dy_full = mappings._gather_along_dim(dy.contiguous(), 0)          # [S, B, H]
        dy2 = dy_full.reshape(-1, wshape[1])
        dw = torch.zeros(wshape, dtype=torch.float32, device=dy.device)
        rc = _lib.lib().nxd_embedding_bwd(_lib.ptr(ids_c), _lib.ptr(dy2), _lib.ptr(dw), _c.c_longlong(ids_c.numel()),
                                          _c.c_int(wshape[1]), _c.c_int(vocab_start), _c.c_int(wshape[0]),
                                          _c.c_int(B), _c.c_int(S), _c.c_int(1), _lib.stream_ptr())
        _lib.check(rc, "nxd_embedding_bwd")
        return None, dw.to(wdtype), None, None


def sp_embedding_push(ids: torch.Tensor, weight: torch.Tensor, vocab_start: int, sym) -> torch.Tensor:
    return _SPEmbeddingPushFn.apply(ids, weight, vocab_start, sym)


def vocab_parallel_embedding(ids: torch.Tensor, weight: torch.Tensor, vocab_start: int,
                             seq_first_out: bool = False) -> torch.Tensor:
    """Rows whose id is outside ``[vocab_start, vocab_start + weight.shape[0])`` come back as zeros
    (the TP all-reduce / reduce-scatter fills them in).  ``seq_first_out`` writes ``[S, B, H]``."""
    if weight.is_cuda and weight.dtype == torch.bfloat16 and weight.shape[1] % 8 == 0 and _lib.available():
        return _VocabParallelEmbeddingFn.apply(ids, weight, vocab_start, seq_first_out)
    local = ids - vocab_start
    mask = (local < 0) | (local >= weight.shape[0])
    out = torch.nn.functional.embedding(local.masked_fill(mask, 0), weight)
    out = out.masked_fill(mask.unsqueeze(-1), 0.0)
    if seq_first_out:
        out = out.transpose(0, 1).contiguous()
    return out
