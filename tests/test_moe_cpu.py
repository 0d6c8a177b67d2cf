"""MoE modules against a plain per-token golden (the reference compares with HF-Mixtral / S-BASE goldens,
``test/unit_test/modules/moe/test_impl_correctness.py``): all three expert paths, both routers, the aux loss, and
expert parallelism (EP=2 over gloo) against EP=1."""
from __future__ import annotations

import math

import pytest
import torch
import torch.nn.functional as F

from conftest import run_distributed


def _golden(x, router_w, gate_up, down, top_k, normalize, capacity=None):
    """x (T,H); gate_up (E,H,2I); down (E,I,H).  Tokens over capacity are dropped in token order."""
    T, H = x.shape
    E = router_w.shape[0]
    logits = x @ router_w.t()
    aff = torch.softmax(logits.float(), dim=-1)
    _, idx = torch.topk(logits, top_k, dim=-1)
    out = torch.zeros_like(x)
    used = [0] * E
    kept = torch.zeros(T, E)
    for t in range(T):
        for e in idx[t].tolist():
            if capacity is None or used[e] < capacity:
                used[e] += 1
                kept[t, e] = 1
    w = aff * kept
    if normalize:
        w = F.normalize(w, p=1.0, dim=1)
    for t in range(T):
        for e in range(E):
            if kept[t, e]:
                gu = x[t] @ gate_up[e]
                g, u = gu.chunk(2)
                out[t] += w[t, e] * ((F.silu(g) * u) @ down[e])
    return out, logits, idx


def _build(E=4, k=2, H=32, I=48, cf=None, normalize=True):
    from neuronx_distributed_llama3_2_b200.modules.moe import MoE, ExpertMLPs, RouterTopK

    torch.manual_seed(0)
    init = lambda w: torch.nn.init.normal_(w, std=0.2)
    router = RouterTopK(E, k, H)
    experts = ExpertMLPs(E, k, H, I, "silu", True, cf, normalize_top_k_affinities=normalize, init_method=init,
                         output_layer_init_method=init)
    return MoE(router, experts, return_router_logits=True, return_expert_index=True)


@pytest.mark.parametrize("cf", [None, 1.0, 0.5])
def test_moe_training_paths_match_golden(cf):
    moe = _build(cf=cf)
    moe.train()
    S, B, H = 12, 2, 32
    x = torch.randn(S, B, H, requires_grad=True)
    out, logits, idx = moe(x)
    T, E, k = S * B, 4, 2
    cap = None if cf is None else min(math.ceil(T * k * cf / E), T)
    gu, down = moe.expert_mlps.mlp_op.gate_up_proj.weight, moe.expert_mlps.mlp_op.down_proj.weight
    ref, ref_logits, ref_idx = _golden(x.detach().reshape(T, H), moe.router.linear_router.weight.detach(), gu.detach(),
                                       down.detach(), k, True, cap)
    torch.testing.assert_close(logits, ref_logits, atol=1e-5, rtol=1e-5)
    assert torch.equal(idx.sort(-1).values, ref_idx.sort(-1).values)
    torch.testing.assert_close(out.reshape(T, H), ref, atol=1e-4, rtol=1e-4)
    out.sum().backward()
    assert x.grad is not None and gu.grad is not None and moe.router.linear_router.weight.grad is not None
    assert getattr(gu, "expert_model_parallel", False)


def test_moe_eval_token_generation_selective_loading():
    moe = _build(E=8, k=2, cf=None)
    moe.eval()
    x = torch.randn(1, 2, 32)            # seq_len 1, batch 2 -> T*k/E = 0.5 < 1 -> selective loading
    out = moe(x)[0]
    gu, down = moe.expert_mlps.mlp_op.gate_up_proj.weight, moe.expert_mlps.mlp_op.down_proj.weight
    ref, _, _ = _golden(x.reshape(2, 32), moe.router.linear_router.weight.detach(), gu.detach(), down.detach(), 2, True)
    torch.testing.assert_close(out.reshape(2, 32), ref, atol=1e-4, rtol=1e-4)
    xl = torch.randn(16, 2, 32)          # context encoding -> dropless all-experts path
    outl = moe(xl)[0]
    refl, _, _ = _golden(xl.reshape(32, 32), moe.router.linear_router.weight.detach(), gu.detach(), down.detach(), 2, True)
    torch.testing.assert_close(outl.reshape(32, 32), refl, atol=1e-4, rtol=1e-4)


def test_router_sinkhorn_and_loss():
    from neuronx_distributed_llama3_2_b200.modules.moe import RouterSinkhorn, load_balancing_loss_func

    torch.manual_seed(0)
    r = RouterSinkhorn(num_experts=4, top_k=1, hidden_size=16)
    x = torch.randn(64, 16)
    r.train()
    _, aff, idx = r(x)
    counts = torch.bincount(idx.reshape(-1), minlength=4)
    assert counts.max() - counts.min() <= 8          # Sinkhorn balances the assignment
    assert aff.min() >= 0 and aff.max() <= 1
    r.eval()
    logits, _, idx_eval = r(x)
    assert torch.equal(idx_eval.reshape(-1), logits.argmax(-1))
    with pytest.raises(NotImplementedError):
        RouterSinkhorn(4, 2, 16)
    # aux loss: perfectly uniform routing gives 1.0
    uni = torch.zeros(128, 4)
    uni[torch.arange(128), torch.arange(128) % 4] = 10.0
    loss = load_balancing_loss_func(uni, 4, 1)
    assert abs(loss.item() - 1.0) < 1e-3
    skew = torch.zeros(128, 4)
    skew[:, 0] = 10.0
    assert load_balancing_loss_func(skew, 4, 1).item() > 3.5
    # Switch-Transformer formula with the reference scaling (loss_function.py:5-48) for top-2
    g = torch.randn(96, 8)
    p = torch.softmax(g, -1)
    sel = torch.topk(p, 2, -1).indices
    mask = F.one_hot(sel, 8).float()
    ref = (mask.mean(0) * p.mean(0).unsqueeze(0)).sum() * 8 / 2     # reference scaling: E / top_k
    assert abs(load_balancing_loss_func(g, 8, 2).item() - ref.item()) < 1e-4


def _ep_worker(rank, world):
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    ps.initialize_model_parallel(tensor_model_parallel_size=1, expert_model_parallel_size=2)
    assert ps.get_expert_model_parallel_size() == 2 and ps.get_data_parallel_size() == 2
    moe = _build(E=4, k=2, cf=1.0)      # each EP rank owns 2 experts
    moe.train()
    full = _build_full_weights()
    with torch.no_grad():
        moe.router.linear_router.weight.copy_(full["router"])
        moe.expert_mlps.mlp_op.gate_up_proj.weight.copy_(full["gate_up"][rank * 2:(rank + 1) * 2])
        moe.expert_mlps.mlp_op.down_proj.weight.copy_(full["down"][rank * 2:(rank + 1) * 2])
    torch.manual_seed(100 + rank)       # every DP/EP rank has its own tokens
    x = torch.randn(8, 2, 32)
    out = moe(x)[0]
    T = 16
    ref, _, _ = _golden(x.reshape(T, 32), full["router"], full["gate_up"], full["down"], 2, True,
                        capacity=min(math.ceil(T * 2 * 1.0 / 4), T))
    torch.testing.assert_close(out.reshape(T, 32), ref, atol=1e-4, rtol=1e-4)
    out.sum().backward()
    assert moe.expert_mlps.mlp_op.down_proj.weight.grad is not None
    # dropless (capacity_factor None) with experts spread over EP ranks: runs as capacity = T, nothing is dropped
    dropless = _build(E=4, k=2, cf=None)
    dropless.train()
    dropless.load_state_dict(moe.state_dict())
    ref2, _, _ = _golden(x.reshape(T, 32), full["router"], full["gate_up"], full["down"], 2, True, capacity=None)
    torch.testing.assert_close(dropless(x)[0].reshape(T, 32), ref2, atol=1e-4, rtol=1e-4)


def _build_full_weights():
    torch.manual_seed(42)
    return {"router": torch.randn(4, 32) * 0.2, "gate_up": torch.randn(4, 32, 96) * 0.2, "down": torch.randn(4, 48, 32) * 0.2}


def test_moe_expert_parallel_matches_golden():
    run_distributed(_ep_worker, 2)
