"""LoRA on the tensor-parallel layers (reference ``src/neuronx_distributed/modules/lora/tp_layer.py:19-207``).

The adapter is sharded the same way as its base layer so no extra collective is needed:
  * RowParallel base  (input split)   A = RowParallelLinear(in/tp -> r, reduces like the base), B = nn.Linear(r -> out)
  * ColumnParallel base (output split) A = nn.Linear(in -> r) replicated,  B = ColumnParallelLinear(r -> out/tp)
  * GQA QKV base                       A shared, B_q / B_k / B_v column-parallel (KV with the base's replication)
"""
from __future__ import annotations

import torch
import torch.nn as nn

from ...parallel.layers import ColumnParallelLinear, RowParallelLinear
from ..qkv_linear import GQAQKVColumnParallelLinear
from .layer import LoraLayer


class LoraParallelLinear(nn.Module, LoraLayer):
    def __init__(self, base_layer, lora_config):
        super().__init__()
        self.init_lora(base_layer, lora_config)
        dt = base_layer.weight.dtype
        dev = base_layer.weight.device
        self.is_row = isinstance(base_layer, RowParallelLinear)
        if self.is_row:
            self.lora_A = RowParallelLinear(base_layer.input_size, self.lora_rank, bias=False,
                                            input_is_parallel=base_layer.input_is_parallel,
                                            sequence_parallel_enabled=base_layer.sequence_parallel_enabled, dtype=dt,
                                            device=dev)
            self.lora_B = nn.Linear(self.lora_rank, base_layer.output_size, bias=False, dtype=dt, device=dev)
            if base_layer.sequence_parallel_enabled:
                setattr(self.lora_B.weight, "sequence_parallel_enabled", True)
        else:
            self.lora_A = nn.Linear(base_layer.input_size, self.lora_rank, bias=False, dtype=dt, device=dev)
            self.lora_B = ColumnParallelLinear(self.lora_rank, base_layer.output_size, bias=False, gather_output=False,
                                               dtype=dt, device=dev)
            self._sp = base_layer.sequence_parallel_enabled
            if self._sp:
                setattr(self.lora_A.weight, "sequence_parallel_enabled", True)
        self.reset_lora_parameters(self.lora_A.weight, self.lora_B.weight)

    def get_delta_weight(self) -> torch.Tensor:
        return (self.lora_B.weight.float() @ self.lora_A.weight.float()) * self.scaling

    def forward(self, x, *a, **kw):
        out = self.base_layer(x, *a, **kw)
        bias = None
        if isinstance(out, tuple):
            out, bias = out
        if not self.merged:
            h = self.lora_dropout(x)
            if self.is_row:
                delta = self.lora_B(self.lora_A(h))
            else:
                if self._sp:  # the base gathers the sequence inside its GEMM; the tiny adapter gathers explicitly
                    from ...parallel import mappings

                    h = mappings.gather_from_sequence_parallel_region(h, to_model_parallel=True)
                delta = self.lora_B(self.lora_A(h))
                if getattr(self.base_layer, "gather_output", False):
                    from ...parallel import mappings

                    delta = mappings.gather_from_tensor_model_parallel_region(delta)
            out = out + delta * self.scaling
        return (out, bias) if bias is not None else out


class LoraGQAQKVParallelLinear(nn.Module, LoraLayer):
    def __init__(self, base_layer: GQAQKVColumnParallelLinear, lora_config):
        super().__init__()
        self.init_lora(base_layer, lora_config)
        w = base_layer.weight_qkv if base_layer.fuse_qkv else base_layer.weight_q
        dt, dev = w.dtype, w.device
        q_out, kv_out = base_layer.output_sizes
        m = base_layer.kv_size_multiplier
        self.lora_A = nn.Linear(base_layer.input_size, self.lora_rank, bias=False, dtype=dt, device=dev)
        self.lora_B_q = ColumnParallelLinear(self.lora_rank, q_out, bias=False, gather_output=False, dtype=dt, device=dev)
        self.lora_B_k = ColumnParallelLinear(self.lora_rank, kv_out * m, bias=False, gather_output=False, dtype=dt,
                                             device=dev)
        self.lora_B_v = ColumnParallelLinear(self.lora_rank, kv_out * m, bias=False, gather_output=False, dtype=dt,
                                             device=dev)
        self._sp = base_layer.sequence_parallel_enabled
        if self._sp:
            setattr(self.lora_A.weight, "sequence_parallel_enabled", True)
        for b in (self.lora_B_q, self.lora_B_k, self.lora_B_v):
            self.reset_lora_parameters(self.lora_A.weight, b.weight)

    def get_qkv(self, layer=None):
        """The three B matrices (reference ``tp_layer.py:148-149`` reads them off a GQA-QKV adapter layer)."""
        layer = layer or self
        if hasattr(layer, "lora_B_q"):
            return layer.lora_B_q.weight, layer.lora_B_k.weight, layer.lora_B_v.weight
        return layer.weight_q, layer.weight_k, layer.weight_v

    def _adapter_pairs(self):
        return [(self.lora_A.weight, b, False) for b in self.get_qkv()]

    def get_delta_weight(self):
        a = self.lora_A.weight.float()
        return tuple((b.float() @ a) * self.scaling for b in self.get_qkv())

    def merge(self) -> None:
        if self.merged:
            return
        dq, dk, dv = self.get_delta_weight()
        b = self.base_layer
        with torch.no_grad():
            if b.fuse_qkv:
                b.weight_qkv.add_(torch.cat([dq, dk, dv], dim=0).to(b.weight_qkv.dtype))
            else:
                b.weight_q.add_(dq.to(b.weight_q.dtype)); b.weight_k.add_(dk.to(b.weight_k.dtype))
                b.weight_v.add_(dv.to(b.weight_v.dtype))
        self.merged = True

    def unmerge(self) -> None:
        if not self.merged:
            return
        dq, dk, dv = self.get_delta_weight()
        b = self.base_layer
        with torch.no_grad():
            if b.fuse_qkv:
                b.weight_qkv.sub_(torch.cat([dq, dk, dv], dim=0).to(b.weight_qkv.dtype))
            else:
                b.weight_q.sub_(dq.to(b.weight_q.dtype)); b.weight_k.sub_(dk.to(b.weight_k.dtype))
                b.weight_v.sub_(dv.to(b.weight_v.dtype))
        self.merged = False

    def forward(self, x, return_fused: bool = False, **kw):
        out = self.base_layer(x, return_fused=return_fused, **kw) if return_fused else self.base_layer(x, **kw)
        if self.merged:
            return out
        h = self.lora_dropout(x)
        if self._sp:
            from ...parallel import mappings

            h = mappings.gather_from_sequence_parallel_region(h, to_model_parallel=True)
        a = self.lora_A(h)
        dq, dk, dv = (b(a) * self.scaling for b in (self.lora_B_q, self.lora_B_k, self.lora_B_v))
        if return_fused:
            return out + torch.cat([dq, dk, dv], dim=-1)
        q, k, v = out
        return q + dq, k + dk, v + dv
