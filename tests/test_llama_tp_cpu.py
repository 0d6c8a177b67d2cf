"""BASELINE.json config #1: tiny Llama (2 layers, hidden 256), TP=2 + SP over gloo on CPU -- loss/grad parity with
the single-process model, ZeRO-1 parity with torch AdamW, sharded checkpoint round trip + directory layout."""
from __future__ import annotations

import os
import tempfile

import torch

from conftest import run_distributed


import contextlib


@contextlib.contextmanager
def _solo():
    """Pretend tp=1 (manual overrides, the reference's `_MPU_*` mocking hooks) to run the un-sharded oracle."""
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    ps.set_tensor_model_parallel_size(1)
    ps.set_tensor_model_parallel_rank(0)
    try:
        yield
    finally:
        ps.set_tensor_model_parallel_size(None)
        ps.set_tensor_model_parallel_rank(None)


def _full_state(seed=0):
    from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM, tiny_llama

    torch.manual_seed(seed)
    cfg = tiny_llama(sequence_parallel_enabled=False)
    with _solo():
        model = LlamaForCausalLM(cfg)
    return cfg, model


def _shard_from_full(full_model, tp_model, tp, rank):
    """Copy the single-process weights into the TP model (shard by partition_dim / stride; fused QKV by part)."""
    from neuronx_distributed_llama3_2_b200.parallel.layers import create_local_weight

    full = dict(full_model.named_parameters())
    with torch.no_grad():
        for name, p in tp_model.named_parameters():
            w = full[name]
            if name.endswith("qkv_proj.weight_qkv"):
                cfg = full_model.config
                q = cfg.num_attention_heads * cfg.head_dim
                kv = cfg.num_key_value_heads * cfg.head_dim
                parts = torch.split(w, [q, kv, kv], dim=0)
                mine = [torch.chunk(x, tp, dim=0)[rank] for x in parts]
                p.copy_(torch.cat(mine, dim=0))
            elif getattr(p, "tensor_model_parallel", False):
                per = w.shape[p.partition_dim] // tp
                p.copy_(create_local_weight(w, p.partition_dim, per, p.partition_stride, rank=rank, world_size=tp))
            else:
                p.copy_(w)


def _tp_worker(rank, world, tmp):
    from neuronx_distributed_llama3_2_b200 import trainer as nxd
    from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM, tiny_llama
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    cfg_full, full = _full_state()
    ids = torch.randint(0, cfg_full.vocab_size, (2, 64), generator=torch.Generator().manual_seed(7))
    with _solo():
        ref_loss = full(ids, labels=ids)
        ref_loss.backward()

    nxd_cfg = nxd.neuronx_distributed_config(tensor_parallel_size=2, sequence_parallel=True,
                                             optimizer_config={"zero_one_enabled": True, "grad_clipping": True,
                                                               "max_grad_norm": 1.0})
    assert ps.get_tensor_model_parallel_size() == 2
    model = nxd.initialize_parallel_model(nxd_cfg, lambda: LlamaForCausalLM(tiny_llama()), device=torch.device("cpu"))
    _shard_from_full(full, model.module, 2, rank)
    loss = model.run_train(ids, labels=ids)
    assert abs(loss.item() - ref_loss.item()) < 1e-4, (loss.item(), ref_loss.item())

    # grads of TP-sharded params == slice of the full grads; SP-replicated (norm) grads need the TP all-reduce
    from neuronx_distributed_llama3_2_b200.parallel.grads import allreduce_sequence_parallel_gradients
    allreduce_sequence_parallel_gradients(list(model.parameters()))
    full_grads = {n: p.grad for n, p in full.named_parameters()}
    checked = 0
    for n, p in model.module.named_parameters():
        g_full = full_grads[n]
        if n.endswith("weight_qkv"):
            c = cfg_full
            q, kv = c.num_attention_heads * c.head_dim, c.num_key_value_heads * c.head_dim
            parts = torch.split(g_full, [q, kv, kv], dim=0)
            want = torch.cat([torch.chunk(x, 2, dim=0)[rank] for x in parts], dim=0)
        elif getattr(p, "tensor_model_parallel", False):
            from neuronx_distributed_llama3_2_b200.parallel.layers import create_local_weight
            want = create_local_weight(g_full, p.partition_dim, g_full.shape[p.partition_dim] // 2,
                                       p.partition_stride, rank=rank, world_size=2)
        else:
            want = g_full
        torch.testing.assert_close(p.grad, want, atol=2e-5, rtol=1e-3, msg=lambda m: f"{n}: {m}")
        checked += 1
    assert checked > 10

    # ZeRO-1 (dp=1 here) AdamW step == torch.optim.AdamW on the full model with the same clipping
    for p in model.parameters():
        p.grad = None
    opt = nxd.initialize_parallel_optimizer(nxd_cfg, torch.optim.AdamW, model.parameters(), lr=1e-2,
                                            weight_decay=0.01, betas=(0.9, 0.95))
    ref_opt = torch.optim.AdamW(full.parameters(), lr=1e-2, weight_decay=0.01, betas=(0.9, 0.95))
    for step in range(3):
        opt.zero_grad()
        ref_opt.zero_grad()
        ids = torch.randint(0, cfg_full.vocab_size, (2, 64), generator=torch.Generator().manual_seed(100 + step))
        l_tp = model.run_train(ids, labels=ids)
        with _solo():
            l_ref = full(ids, labels=ids)
            l_ref.backward()
        assert abs(l_tp.item() - l_ref.item()) < 5e-4, (step, l_tp.item(), l_ref.item())
        ref_norm = torch.nn.utils.clip_grad_norm_(full.parameters(), 1.0)
        opt.step()
        ref_opt.step()
        assert abs(float(opt.grad_norm) - float(ref_norm)) < 1e-3 * max(1.0, float(ref_norm)), (
            float(opt.grad_norm), float(ref_norm))
    w_tp = dict(model.module.named_parameters())["model.norm.weight"]
    torch.testing.assert_close(w_tp, dict(full.named_parameters())["model.norm.weight"], atol=1e-4, rtol=1e-3)

    # sharded checkpoint: layout + bit-exact round trip (model and optimizer), xser and plain
    for use_xser in (False, True):
        d = os.path.join(tmp, f"ckpt_x{int(use_xser)}")
        nxd.save_checkpoint(d, "step_3", model=model, optimizer=opt, user_content={"step": 3}, use_xser=use_xser)
        shard = f"dp_rank_00_tp_rank_{rank:02d}_pp_rank_00.pt"
        assert os.path.exists(os.path.join(d, "step_3", "done"))
        assert os.path.exists(os.path.join(d, "step_3", "checkpoint"))
        assert os.path.exists(os.path.join(d, "step_3", "model", shard))
        assert os.path.exists(os.path.join(d, "step_3", "optim", shard))
        if use_xser:
            assert os.path.isdir(os.path.join(d, "step_3", "model", shard + ".tensors"))
            assert os.path.exists(os.path.join(d, "step_3", "model", shard + ".info.pt"))
        before = {k: v.clone() for k, v in model.state_dict().items()}
        opt_before = opt.state_dict()
        with torch.no_grad():
            for p in model.parameters():
                p.add_(1.0)
        uc = nxd.load_checkpoint(d, tag=None, model=model, optimizer=opt)
        assert uc == {"step": 3}
        for k, v in model.state_dict().items():
            assert torch.equal(v, before[k]), k
        after = opt.state_dict()
        for i in opt_before["base_state"]:
            assert torch.equal(opt_before["base_state"][i], after["base_state"][i])
            assert torch.equal(opt_before["state"][i]["exp_avg"], after["state"][i]["exp_avg"])
    assert nxd.has_checkpoint(os.path.join(tmp, "ckpt_x0"))


def test_tiny_llama_tp2_sp_zero1_checkpoint():
    with tempfile.TemporaryDirectory() as tmp:
        run_distributed(_tp_worker, 2, tmp)


def test_single_process_defaults():
    """Without torch.distributed everything behaves as tp=pp=dp=1."""
    from neuronx_distributed_llama3_2_b200 import trainer as nxd
    from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM, tiny_llama

    cfg = nxd.neuronx_distributed_config(optimizer_config={"zero_one_enabled": True, "grad_clipping": True,
                                                           "max_grad_norm": 1.0})
    assert cfg["mixed_precision_config"]["use_master_weights"] is True
    torch.manual_seed(0)
    model = nxd.initialize_parallel_model(cfg, lambda: LlamaForCausalLM(tiny_llama()), device=torch.device("cpu"))
    opt = nxd.initialize_parallel_optimizer(cfg, torch.optim.AdamW, model.parameters(), lr=1e-2)
    ids = torch.randint(0, 512, (2, 32))
    losses = []
    for _ in range(8):
        opt.zero_grad()
        losses.append(model.run_train(ids, labels=ids).item())
        opt.step()
    assert losses[-1] < losses[0] - 0.5, losses
