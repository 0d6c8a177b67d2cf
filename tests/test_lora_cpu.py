"""LoRA: layer replacement (plain + tensor-parallel layers), zero-delta at init, merge/unmerge, adapter-only state dict,
save/load (reference ``test/unit_test/modules/lora/test_{lora_layers,lora_models,save_load}.py``)."""
from __future__ import annotations

import os
import tempfile

import torch
import torch.nn as nn

from neuronx_distributed_llama3_2_b200.modules.lora import (LoraConfig, LoraEmbedding, LoraGQAQKVParallelLinear,
                                                            LoraLinear, LoraModel, LoraParallelLinear, get_lora_model)


class _Tiny(nn.Module):
    def __init__(self):
        super().__init__()
        self.emb = nn.Embedding(20, 16)
        self.q_proj = nn.Linear(16, 16)
        self.v_proj = nn.Linear(16, 16)
        self.out = nn.Linear(16, 4)

    def forward(self, ids):
        h = self.emb(ids)
        return self.out(self.q_proj(h) + self.v_proj(h))


def test_lora_plain_layers_merge_and_state_dict():
    torch.manual_seed(0)
    base = _Tiny()
    plain_state = {k: v.clone() for k, v in base.state_dict().items()}
    ids = torch.randint(0, 20, (3, 5))
    ref = base(ids)
    cfg = LoraConfig(enable_lora=True, lora_rank=4, lora_alpha=8, target_modules=["q_proj", "v_proj", "emb"])
    model = get_lora_model(base, cfg)
    assert isinstance(model, LoraModel)
    assert isinstance(base.q_proj, LoraLinear) and isinstance(base.emb, LoraEmbedding) and isinstance(base.out, nn.Linear)
    torch.testing.assert_close(model(ids), ref)            # B = 0 at init: output unchanged
    trainable = [n for n, p in model.named_parameters() if p.requires_grad]
    assert trainable and all("lora_" in n for n in trainable)
    with torch.no_grad():
        base.q_proj.lora_B.weight.normal_()
        base.emb.lora_embedding_A.normal_()
    out = model(ids)
    assert not torch.allclose(out, ref)
    model.merge_lora()
    torch.testing.assert_close(model(ids), out, atol=1e-5, rtol=1e-5)
    model.unmerge_lora()
    torch.testing.assert_close(model(ids), out, atol=1e-5, rtol=1e-5)
    sd = model.state_dict()
    assert "lora_config" in sd and all("lora_" in k for k in sd if k != "lora_config")
    with tempfile.TemporaryDirectory() as d:
        cfg.lora_save_dir = d
        path = model.save_lora()
        assert os.path.exists(path)
        fresh = _Tiny()
        fresh.load_state_dict(plain_state)
        m2 = LoraModel(fresh, LoraConfig(enable_lora=True, load_lora_from_ckpt=True, lora_save_dir=d))
        torch.testing.assert_close(m2(ids), out, atol=1e-5, rtol=1e-5)


def test_lora_on_tensor_parallel_llama():
    from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM, tiny_llama

    torch.manual_seed(0)
    m = LlamaForCausalLM(tiny_llama())
    ids = torch.randint(0, 512, (2, 16))
    ref = m(ids, labels=ids)
    model = LoraModel(m, LoraConfig(enable_lora=True, lora_rank=8, target_modules=["qkv_proj", "o_proj", "gate_up_proj"]))
    layer0 = m.model.layers[0]
    assert isinstance(layer0.self_attn.qkv_proj, LoraGQAQKVParallelLinear)
    assert isinstance(layer0.self_attn.o_proj, LoraParallelLinear) and layer0.self_attn.o_proj.is_row
    assert isinstance(layer0.mlp.gate_up_proj, LoraParallelLinear) and not layer0.mlp.gate_up_proj.is_row
    loss = model(ids, labels=ids)
    torch.testing.assert_close(loss, ref)
    loss.backward()
    grads = {n for n, p in model.named_parameters() if p.grad is not None}
    assert grads and all("lora_" in n for n in grads)
    with torch.no_grad():
        for mod in m.modules():
            if isinstance(mod, LoraGQAQKVParallelLinear):
                mod.lora_B_q.weight.normal_(std=0.05)
                mod.lora_B_v.weight.normal_(std=0.05)
    changed = model(ids, labels=ids)
    assert abs(changed.item() - ref.item()) > 1e-6
    model.merge_lora()
    torch.testing.assert_close(model(ids, labels=ids), changed, atol=1e-4, rtol=1e-4)
    model.unmerge_lora()
    assert "trainable params" in model.print_trainable_parameters()


def test_lora_model_helpers_base_checkpoint_keys_config_and_two_step_load():
    import json

    torch.manual_seed(0)
    base = _Tiny()
    plain_state = {k: v.clone() for k, v in base.state_dict().items()}
    cfg = LoraConfig(enable_lora=True, lora_rank=2, lora_alpha=4, target_modules=["q_proj", "v_proj"])
    model = get_lora_model(base, cfg)
    t, n = model.get_nb_trainable_parameters()
    assert 0 < t < n and f"{t:,d}" in model.print_trainable_parameters()
    # a BASE checkpoint loads into the wrapped model after its keys are re-pointed at ``base_layer``
    mapped = model.update_state_dict_keys(plain_state)
    assert "q_proj.base_layer.weight" in mapped and "out.weight" in mapped and "q_proj.weight" not in mapped
    missing = model.get_base_model().load_state_dict(mapped, strict=False).missing_keys
    assert missing and all("lora_" in k for k in missing)
    assert any("base_layer" in k for k in model.module_state_dict())
    with torch.no_grad():
        base.v_proj.lora_B.weight.normal_()
    ids = torch.randint(0, 20, (2, 4))
    want = model(ids)
    with tempfile.TemporaryDirectory() as d:
        cfg.lora_save_dir = d
        model.save_lora()
        cfg_path = model.save_config()
        with open(cfg_path) as f:
            saved = json.load(f)
        assert saved["lora_rank"] == 2 and "target_modules" in saved
        fresh = _Tiny()
        fresh.load_state_dict(plain_state)
        m2 = LoraModel(fresh, LoraConfig(enable_lora=True, lora_rank=2, lora_alpha=4, target_modules=["q_proj", "v_proj"],
                                         lora_save_dir=d))
        assert not torch.allclose(m2(ids), want)
        sd = m2.load_checkpoint()                    # step 1: read the file, adopt its configuration
        assert "lora_config" in sd
        m2.load_lora_adapter()                       # step 2: apply the adapter weights
        torch.testing.assert_close(m2(ids), want, atol=1e-6, rtol=1e-6)


def test_lora_layer_update_init_and_model_properties():
    torch.manual_seed(0)
    base = _Tiny()
    cfg = LoraConfig(enable_lora=True, lora_rank=4, lora_alpha=8, target_modules=["q_proj"])
    model = get_lora_model(base, cfg)
    layer = base.q_proj
    ids = torch.randint(0, 20, (2, 3))
    ref = model(ids)
    with torch.no_grad():
        layer.lora_B.weight.normal_()
    assert not torch.allclose(model(ids), ref)
    layer.init_lora_parameters("gaussian")                   # B back to zero -> output back to the base model
    torch.testing.assert_close(model(ids), ref)
    assert float(layer.lora_A.weight.std()) < 0.5            # N(0, 1/r)
    layer.merge()
    layer.update_layer(LoraConfig(enable_lora=True, lora_rank=2, lora_alpha=2, target_modules=["q_proj"]))
    assert layer.lora_A.weight.shape == (2, 16) and not layer.merged and layer.scaling == 1.0
    torch.testing.assert_close(model(ids), ref)              # unmerged before the rebuild: base weight intact
    assert torch.equal(layer.transpose(layer.lora_A.weight), layer.lora_A.weight)
    assert model.dtype == torch.float32 and model.config is None
    model.print_model_info()
