"""The reference's ZeRO-1 sweep (``test/integration/zero1/test_zero1.py:157-176``) over gloo: {TP2, PP2} x model dtype x
optimizer dtype x clipping x max_norm, five steps each with ``grad = param / 100``, compared with an UNSHARDED optimizer
that does the same arithmetic on full tensors (the reference's ``RefOptimizer``, ``zero1/utils.py:91-262``).  World = 4,
so every case also has a real data-parallel pair: the reduce-scatter / sharded AdamW / all-gather path is what runs."""
from __future__ import annotations

import itertools

import torch

from conftest import run_distributed

LR, STEPS = 1e-2, 5


def _cases():
    for mdt, odt, clip, max_norm in itertools.product([torch.float32, torch.bfloat16], [torch.float32, torch.bfloat16],
                                                      [True, False], [1.0, 0.5, 0.2, 2.0]):
        if mdt == torch.float32 and odt == torch.bfloat16:      # the reference skips fp32 weights + bf16 states
            continue
        if not clip and max_norm != 1.0:                        # max_norm has no effect without clipping
            continue
        yield mdt, odt, clip, max_norm


def _make_params(mdt, tp_rank, pp_rank):
    """Two tensor-parallel shards and one replicated tensor per pipeline stage; magnitudes large enough that the
    gradient ``p / 100`` has a norm well above every max_norm in the sweep."""
    from neuronx_distributed_llama3_2_b200.parallel.utils import set_tensor_model_parallel_attributes

    g = torch.Generator().manual_seed(100 + pp_rank)
    full_a = torch.randn(16, 24, generator=g) * 4
    full_b = torch.randn(24, 16, generator=g) * 4
    rep = torch.randn(24, generator=g) * 4
    tp = 2 if tp_rank is not None else 1
    r = tp_rank or 0
    a = full_a.chunk(tp, 0)[r].clone().to(mdt).requires_grad_()
    b = full_b.chunk(tp, 1)[r].clone().to(mdt).requires_grad_()
    c = rep.clone().to(mdt).requires_grad_()
    if tp > 1:
        set_tensor_model_parallel_attributes(a, True, 0, 1)
        set_tensor_model_parallel_attributes(b, True, 1, 1)
    return [a, b, c]


def _reference_run(params, mdt, odt, clip, max_norm, norm_groups, tp):
    """Unsharded oracle: masters / moments in the optimizer dtype, the global norm counts replicated tensors once."""
    import torch.distributed as dist

    master = [p.detach().clone().to(odt) for p in params]
    m = [torch.zeros_like(x) for x in master]
    v = [torch.zeros_like(x) for x in master]
    beta1, beta2, eps, wd = 0.9, 0.999, 1e-8, 1e-2
    out = []
    for step in range(1, STEPS + 1):
        grads = [(p.detach().clone() / 100).to(torch.float32) for p in params]
        if clip:
            sq = torch.zeros((), dtype=torch.float64)
            for p, g in zip(params, grads):
                dup = not getattr(p, "tensor_model_parallel", False)
                sq += g.double().pow(2).sum() / (tp if dup else 1)
            for grp in norm_groups:
                dist.all_reduce(sq, group=grp)
            coef = min(1.0, max_norm / (float(sq.sqrt()) + 1e-6))
            grads = [g * coef for g in grads]
        for i, g in enumerate(grads):
            w = master[i].float()
            w = w * (1 - LR * wd)
            mi = m[i].float() * beta1 + g * (1 - beta1)
            vi = v[i].float() * beta2 + g * g * (1 - beta2)
            m[i], v[i] = mi.to(odt), vi.to(odt)
            w = w - LR * (mi / (1 - beta1 ** step)) / ((vi / (1 - beta2 ** step)).sqrt() + eps)
            master[i] = w.to(odt)
            params[i].data.copy_(master[i].to(mdt))
        out.append(torch.stack([p.detach().double().sum() for p in params]).mean())
    return out


def _zero1_run(params, mdt, odt, clip, max_norm):
    from neuronx_distributed_llama3_2_b200.optimizer import NeuronZero1Optimizer

    opt = NeuronZero1Optimizer(params, torch.optim.AdamW, optimizer_dtype=odt, grad_clipping=clip, max_norm=max_norm,
                               lr=LR, weight_decay=1e-2)
    out = []
    for _ in range(STEPS):
        for p in params:
            p.grad = p.detach().clone() / 100
        opt.step()
        opt.zero_grad()
        out.append(torch.stack([p.detach().double().sum() for p in params]).mean())
    return out, opt


def _worker(rank, world, layout):
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    tp, pp = (2, 1) if layout == "tp" else (1, 2)
    ps.initialize_model_parallel(tensor_model_parallel_size=tp, pipeline_model_parallel_size=pp)
    assert ps.get_data_parallel_size() == 2
    tp_rank = ps.get_tensor_model_parallel_rank() if tp > 1 else None
    pp_rank = ps.get_pipeline_model_parallel_rank()
    groups = []
    if tp > 1:
        groups.append(ps.get_tensor_model_parallel_group())
    if pp > 1:
        groups.append(ps.get_pipeline_model_parallel_group())
    n = 0
    for mdt, odt, clip, max_norm in _cases():
        got, opt = _zero1_run(_make_params(mdt, tp_rank, pp_rank), mdt, odt, clip, max_norm)
        assert opt.buckets[0].dp == 2
        want = _reference_run(_make_params(mdt, tp_rank, pp_rank), mdt, odt, clip, max_norm, groups, tp)
        tol = 5e-2 if torch.bfloat16 in (mdt, odt) else 1e-4    # the reference's tolerance; fp32 / fp32 must be tight
        for s, (a, b) in enumerate(zip(got, want)):
            torch.testing.assert_close(a, b, rtol=tol, atol=tol, msg=lambda m: f"{layout} {mdt} {odt} clip={clip} "
                                       f"max_norm={max_norm} step {s}: {m}")
        n += 1
    assert n == 15


def test_zero1_matrix_tp2_dp2():
    run_distributed(_worker, 4, "tp", timeout=300.0)


def test_zero1_matrix_pp2_dp2():
    run_distributed(_worker, 4, "pp", timeout=300.0)
