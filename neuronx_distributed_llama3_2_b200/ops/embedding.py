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
