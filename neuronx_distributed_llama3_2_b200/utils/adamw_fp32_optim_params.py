"""``AdamW_FP32OptimParams``: AdamW that keeps fp32 master copies and fp32 moments for low-precision
parameters (reference ``src/neuronx_distributed/utils/adamw_fp32_optim_params.py:31-155``, which
also plays an fp64 trick under ``XLA_DOWNCAST_BF16`` that has no meaning off XLA).

Update rule (HF-style, as the reference ``:132-153``): bias-corrected Adam step, then decoupled weight decay
``p -= lr * wd * p``.  On CUDA the whole update of a parameter is ONE launch of the fused sm_100a kernel
(``csrc/optimizer.cu``), which also writes the rounded low-precision parameter.
"""
from __future__ import annotations

import math

import torch

from ..ops import optim as optim_ops


class AdamW_FP32OptimParams(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.0, correct_bias=True,
                 no_deprecation_warning: bool = True, capturable: bool = False, differentiable: bool = False):
        # the last three exist for signature parity with the HF-derived reference optimizer; nothing here depends on them
        if lr < 0.0:
            raise ValueError(f"Invalid learning rate: {lr}")
        if not 0.0 <= betas[0] < 1.0 or not 0.0 <= betas[1] < 1.0:
            raise ValueError(f"Invalid beta parameters: {betas}")
        if eps < 0.0:
            raise ValueError(f"Invalid epsilon value: {eps}")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay,
                                      correct_bias=correct_bias))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group in self.param_groups:
            beta1, beta2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                if p.grad.is_sparse:
                    raise RuntimeError("AdamW_FP32OptimParams does not support sparse gradients")
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p, dtype=torch.float32)
                    st["exp_avg_sq"] = torch.zeros_like(p, dtype=torch.float32)
                    st["master"] = p.detach().clone().float() if p.dtype != torch.float32 else None
                st["step"] += 1
                master = st["master"] if st["master"] is not None else p.data
                fused = (p.is_cuda and group["correct_bias"] and p.is_contiguous() and p.grad.is_contiguous()
                         and p.grad.dtype in (torch.float32, torch.bfloat16))
                if fused:
                    optim_ops.adamw_step_(master.view(-1), st["exp_avg"].view(-1), st["exp_avg_sq"].view(-1),
                                          p.grad.view(-1), p.data.view(-1) if st["master"] is not None else None,
                                          group["lr"], beta1, beta2, group["eps"], group["weight_decay"], st["step"])
                    continue
                g = p.grad.float()
                st["exp_avg"].mul_(beta1).add_(g, alpha=1.0 - beta1)
                st["exp_avg_sq"].mul_(beta2).addcmul_(g, g, value=1.0 - beta2)
                step_size = group["lr"]
                denom = st["exp_avg_sq"].sqrt().add_(group["eps"])
                if group["correct_bias"]:
                    bc1 = 1.0 - beta1 ** st["step"]
                    bc2 = 1.0 - beta2 ** st["step"]
                    step_size = step_size * math.sqrt(bc2) / bc1
                master.addcdiv_(st["exp_avg"], denom, value=-step_size)
                if group["weight_decay"] > 0.0:
                    master.add_(master, alpha=-group["lr"] * group["weight_decay"])
                if st["master"] is not None:
                    p.data.copy_(master.to(p.dtype))
        return loss
