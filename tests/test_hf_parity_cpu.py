"""Every model family against the Hugging Face ``transformers`` implementation of the same architecture on CPU, through
the checkpoint-name converters (the reference's accuracy check compares against HF on CPU, ``examples/inference/
runner.py:313-409``).  Random weights, tiny shapes, fp32: logits must agree to float rounding."""
from __future__ import annotations

import pytest
import torch

transformers = pytest.importorskip("transformers")

from neuronx_distributed_llama3_2_b200 import models as M  # noqa: E402


def _ids(vocab=500, B=2, T=12):
    return torch.randint(1, vocab, (B, T), generator=torch.Generator().manual_seed(1))


def test_llama_matches_hf_including_llama3_rope_scaling():
    from transformers import LlamaConfig, LlamaForCausalLM

    from neuronx_distributed_llama3_2_b200.scripts.checkpoint_converter import CheckpointConverterBase

    torch.manual_seed(0)
    c = M.tiny_llama(sequence_parallel_enabled=False, max_position_embeddings=16384)
    mine = M.LlamaForCausalLM(c).eval()
    conv = CheckpointConverterBase(c.num_hidden_layers, c.num_attention_heads, c.num_key_value_heads, c.head_dim,
                                   tie_word_embeddings=True)
    hf = LlamaForCausalLM(LlamaConfig(
        vocab_size=c.vocab_size, hidden_size=c.hidden_size, intermediate_size=c.intermediate_size, num_hidden_layers=2,
        num_attention_heads=8, num_key_value_heads=2, head_dim=32, rms_norm_eps=c.rms_norm_eps, rope_theta=c.rope_theta,
        rope_scaling=dict(c.rope_scaling), max_position_embeddings=16384, tie_word_embeddings=True,
        attn_implementation="eager")).eval()
    res = hf.load_state_dict(conv.nxd_full_to_hf(mine.state_dict()), strict=False)
    assert not res.unexpected_keys and set(res.missing_keys) <= {"lm_head.weight"}
    ids = _ids()
    with torch.no_grad():
        torch.testing.assert_close(mine(ids).transpose(0, 1), hf(ids).logits, atol=2e-5, rtol=1e-4)
    # and the other direction: HF names -> fused QKV / gate_up
    back = conv.hf_to_nxd_full({k: v for k, v in hf.state_dict().items()})
    assert all(torch.equal(back[k], v) for k, v in mine.state_dict().items())


def test_mixtral_matches_hf_through_checkpoint_names():
    from transformers import MixtralConfig, MixtralForCausalLM

    torch.manual_seed(0)
    c = M.tiny_mixtral(sequence_parallel_enabled=False, capacity_factor=None)      # dropless, as HF
    mine = M.MixtralForCausalLM(c).eval()
    disk = M.convert_mixtral_state_dict_to_hf(mine.state_dict(), c)                 # classic on-disk names (w1 / w2 / w3)
    back = M.convert_mixtral_hf_state_dict(disk, c)
    assert all(torch.equal(back[k], v) for k, v in mine.state_dict().items())
    hf = MixtralForCausalLM(MixtralConfig(
        vocab_size=c.vocab_size, hidden_size=c.hidden_size, intermediate_size=c.intermediate_size, num_hidden_layers=2,
        num_attention_heads=4, num_key_value_heads=2, head_dim=32, num_local_experts=4, num_experts_per_tok=2,
        rope_theta=c.rope_theta, rms_norm_eps=c.rms_norm_eps, max_position_embeddings=512, tie_word_embeddings=False,
        sliding_window=None, attn_implementation="eager")).eval()
    names = set(hf.state_dict())
    if "model.layers.0.mlp.experts.gate_up_proj" in names:        # transformers >= 5 keeps the experts stacked in memory
        E, sd = c.num_local_experts, {k: v for k, v in disk.items() if "block_sparse_moe" not in k}
        for i in range(c.num_hidden_layers):
            p, moe = f"model.layers.{i}.mlp.", f"model.layers.{i}.block_sparse_moe."
            sd[p + "gate.weight"] = disk[moe + "gate.weight"]
            sd[p + "experts.gate_up_proj"] = torch.stack(
                [torch.cat([disk[moe + f"experts.{e}.w1.weight"], disk[moe + f"experts.{e}.w3.weight"]], 0) for e in range(E)])
            sd[p + "experts.down_proj"] = torch.stack([disk[moe + f"experts.{e}.w2.weight"] for e in range(E)])
    else:
        sd = disk
    res = hf.load_state_dict(sd, strict=False)
    assert not res.unexpected_keys and not res.missing_keys, res
    ids = _ids()
    with torch.no_grad():
        torch.testing.assert_close(mine(ids).transpose(0, 1), hf(ids).logits, atol=2e-5, rtol=1e-4)


def test_gpt_neox_matches_hf():
    from transformers import GPTNeoXConfig, GPTNeoXForCausalLM

    torch.manual_seed(0)
    c = M.tiny_gpt_neox()
    mine = M.GPTNeoXForCausalLM(c).eval()
    hf = GPTNeoXForCausalLM(GPTNeoXConfig(
        vocab_size=c.vocab_size, hidden_size=c.hidden_size, num_hidden_layers=2, num_attention_heads=4,
        intermediate_size=c.intermediate_size, rotary_pct=c.rotary_pct, rotary_emb_base=c.rotary_emb_base,
        use_parallel_residual=True, layer_norm_eps=c.layer_norm_eps, max_position_embeddings=c.max_position_embeddings,
        hidden_act="gelu", attention_bias=True, tie_word_embeddings=False, attn_implementation="eager")).eval()
    sd = {(k if k.startswith("embed_out") else "gpt_neox." + k): v for k, v in mine.state_dict().items()}
    res = hf.load_state_dict(sd, strict=False)
    assert not res.unexpected_keys and not res.missing_keys, res
    ids = _ids()
    with torch.no_grad():
        torch.testing.assert_close(mine(ids).transpose(0, 1), hf(ids).logits, atol=2e-5, rtol=1e-4)


def test_bert_matches_hf_with_and_without_padding_mask():
    from transformers import BertConfig, BertForPreTraining

    torch.manual_seed(0)
    c = M.tiny_bert()                                              # vocab 512: the HF vocab of 500 gets padded
    hf = BertForPreTraining(BertConfig(
        vocab_size=500, hidden_size=c.hidden_size, num_hidden_layers=2, num_attention_heads=4,
        intermediate_size=c.intermediate_size, max_position_embeddings=c.max_position_embeddings, type_vocab_size=2,
        layer_norm_eps=c.layer_norm_eps, hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0,
        attn_implementation="eager")).eval()
    with torch.no_grad():
        hf.cls.predictions.bias.normal_()
    mine = M.BertForPreTraining(c).eval()
    res = mine.load_state_dict(M.convert_bert_hf_state_dict(hf.state_dict(), c), strict=False)
    assert not res.unexpected_keys and not res.missing_keys, res
    ids = _ids()
    tt = torch.randint(0, 2, ids.shape, generator=torch.Generator().manual_seed(2))
    mask = torch.ones_like(ids)
    mask[1, 9:] = 0
    with torch.no_grad():
        for am in (None, mask):
            ref = hf(ids, token_type_ids=tt, attention_mask=am)
            logits, nsp = mine(ids, tt, am)
            got = logits.transpose(0, 1)
            keep = torch.ones_like(ids, dtype=torch.bool) if am is None else am.bool()
            torch.testing.assert_close(got[..., :500][keep], ref.prediction_logits[keep], atol=2e-5, rtol=1e-4)
            torch.testing.assert_close(nsp, ref.seq_relationship_logits, atol=2e-5, rtol=1e-4)
            assert bool((got[..., 500:] < -1e3).all())             # padded vocabulary rows can never be predicted


def test_dbrx_matches_hf():
    """HF's in-memory expert weights are ``[E, d_model, ffn]`` slabs in transformers 5 (``x @ w1``), the published
    checkpoints -- which ``convert_dbrx_hf_state_dict`` reads -- hold ``[E * ffn, d_model]`` (``x @ w1.T``): the slabs are
    transposed here before conversion.  That HF version also sizes the router by ``ffn_hidden_size`` and ignores
    ``attn_config.rope_theta``, hence ffn == d_model and theta 10000 in this comparison."""
    from transformers import DbrxConfig, DbrxForCausalLM
    from transformers.models.dbrx.configuration_dbrx import DbrxAttentionConfig, DbrxFFNConfig

    torch.manual_seed(0)
    c = M.tiny_dbrx(sequence_parallel_enabled=False, capacity_factor=None, intermediate_size=128, rope_theta=10000.0)
    hf = DbrxForCausalLM(DbrxConfig(
        d_model=c.hidden_size, n_heads=4, n_layers=2, max_seq_len=512, vocab_size=c.vocab_size,
        attn_config=DbrxAttentionConfig(kv_n_heads=2, clip_qkv=c.clip_qkv, rope_theta=c.rope_theta),
        ffn_config=DbrxFFNConfig(ffn_hidden_size=128, moe_num_experts=4, moe_top_k=2, moe_normalize_expert_weights=1.0,
                                 hidden_size=c.hidden_size), attn_implementation="eager")).eval()
    if abs(float(hf.transformer.rotary_emb.inv_freq[1]) - 10000.0 ** (-2 / 32)) > 1e-4:
        pytest.skip("this transformers version derives the DBRX rotary base differently")
    with torch.no_grad():
        for p in hf.parameters():
            if p.dim() > 1:
                p.normal_(std=0.08)                                # large enough for the QKV clamp (0.5) to bite
    sd = dict(hf.state_dict())
    for k in list(sd):
        if k.endswith((".w1", ".v1", ".w2")):
            sd[k] = sd[k].reshape(4, 128, 128).transpose(1, 2).reshape(512, 128).contiguous()
    mine = M.DbrxForCausalLM(c).eval()
    res = mine.load_state_dict(M.convert_dbrx_hf_state_dict(sd, c), strict=False)
    assert not res.unexpected_keys and not res.missing_keys, res
    ids = _ids()
    with torch.no_grad():
        torch.testing.assert_close(mine(ids).transpose(0, 1), hf(ids).logits, atol=5e-5, rtol=1e-4)
