"""HF-named full checkpoint <-> (tp, pp)-sharded NxD checkpoint round trips, incl. KV-head replication and the match with
what the parallel model itself would hold (reference ``scripts/checkpoint_converter.py`` behaviour)."""
from __future__ import annotations

import os
import tempfile

import torch

from neuronx_distributed_llama3_2_b200.scripts.checkpoint_converter import CheckpointConverterBase, q_head_permutation


def _hf_state(L=4, H=64, nq=8, nkv=2, D=8, I=96, V=50, tied=True):
    torch.manual_seed(0)
    sd = {"model.embed_tokens.weight": torch.randn(V, H), "model.norm.weight": torch.randn(H)}
    for i in range(L):
        p = f"model.layers.{i}."
        sd[p + "self_attn.q_proj.weight"] = torch.randn(nq * D, H)
        sd[p + "self_attn.k_proj.weight"] = torch.randn(nkv * D, H)
        sd[p + "self_attn.v_proj.weight"] = torch.randn(nkv * D, H)
        sd[p + "self_attn.o_proj.weight"] = torch.randn(H, nq * D)
        sd[p + "mlp.gate_proj.weight"] = torch.randn(I, H)
        sd[p + "mlp.up_proj.weight"] = torch.randn(I, H)
        sd[p + "mlp.down_proj.weight"] = torch.randn(H, I)
        sd[p + "input_layernorm.weight"] = torch.randn(H)
        sd[p + "post_attention_layernorm.weight"] = torch.randn(H)
    if not tied:
        sd["lm_head.weight"] = torch.randn(V, H)
    return sd


def test_round_trip_tp2_pp2_xser_and_plain():
    hf = _hf_state()
    for use_xser in (False, True):
        conv = CheckpointConverterBase(4, 8, 2, 8, tp_size=2, pp_size=2, tie_word_embeddings=True)
        with tempfile.TemporaryDirectory() as d:
            conv.convert_from_full_state(hf, d, use_xser=use_xser)
            for t in range(2):
                for s in range(2):
                    assert os.path.exists(os.path.join(d, "model", f"dp_rank_00_tp_rank_{t:02d}_pp_rank_{s:02d}.pt"))
            first = torch.load(os.path.join(d, "model", "dp_rank_00_tp_rank_00_pp_rank_00.pt"), weights_only=False) \
                if not use_xser else None
            if first is not None:
                assert "model.embed_tokens.weight" in first and "model.layers.0.mlp.gate_up_proj.weight" in first
                assert "model.layers.3.mlp.down_proj.weight" not in first and "model.norm.weight" not in first
                assert first["model.layers.0.self_attn.qkv_proj.weight_qkv"].shape == ((8 + 2 + 2) * 8 // 2, 64)
            back = conv.convert_to_full_state(d)
        assert set(back) == set(hf)
        for k in hf:
            assert torch.equal(back[k], hf[k]), k


def test_kv_replication_permutation_and_round_trip():
    # 2 KV heads, tp = 4 -> every KV head is replicated twice; rank r holds KV head r % 2
    perm = q_head_permutation(n_q=8, n_kv=2, tp=4, kv_mult=2)
    assert perm == [0, 1, 4, 5, 2, 3, 6, 7]      # rank0: q0,q1 (kv0)  rank1: q4,q5 (kv1)  rank2: q2,q3 (kv0)  rank3: q6,q7
    hf = _hf_state(tied=False)
    conv = CheckpointConverterBase(4, 8, 2, 8, tp_size=4, pp_size=1, kv_size_multiplier=2, tie_word_embeddings=False)
    full = conv.hf_to_nxd_full(hf)
    r1 = conv.shard_tp(full, 1)
    qkv = r1["model.layers.0.self_attn.qkv_proj.weight_qkv"]
    assert qkv.shape[0] == (2 + 1 + 1) * 8
    torch.testing.assert_close(qkv[:16], hf["model.layers.0.self_attn.q_proj.weight"][4 * 8:6 * 8])   # q heads 4, 5
    torch.testing.assert_close(qkv[16:24], hf["model.layers.0.self_attn.k_proj.weight"][8:16])        # kv head 1
    with tempfile.TemporaryDirectory() as d:
        conv.convert_from_full_state(hf, d)
        back = conv.convert_to_full_state(d)
    for k in hf:
        assert torch.equal(back[k], hf[k]), k


def test_converted_shards_load_into_parallel_model():
    """A converted tp=1 checkpoint loads into our Llama and reproduces a plain-PyTorch Llama forward on the HF weights."""
    from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM, tiny_llama

    cfg = tiny_llama(sequence_parallel_enabled=False)
    hf = _hf_state(L=cfg.num_hidden_layers, H=cfg.hidden_size, nq=cfg.num_attention_heads, nkv=cfg.num_key_value_heads,
                   D=cfg.head_dim, I=cfg.intermediate_size, V=cfg.vocab_size)
    hf = {k: v * 0.05 if v.dim() == 2 else 1 + 0.1 * v for k, v in hf.items()}
    conv = CheckpointConverterBase(cfg.num_hidden_layers, cfg.num_attention_heads, cfg.num_key_value_heads, cfg.head_dim)
    model = LlamaForCausalLM(cfg)
    model.load_state_dict(conv.shard_tp(conv.hf_to_nxd_full(hf), 0))
    ids = torch.randint(0, cfg.vocab_size, (1, 10))
    logits = model(ids)[:, 0]                                   # [S, V]

    def rms(x, w):
        return x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + cfg.rms_norm_eps) * w

    from neuronx_distributed_llama3_2_b200.ops.rope import build_rope_cache

    cos, sin = build_rope_cache(10, cfg.head_dim, cfg.rope_theta, cfg.rope_scaling)

    def rope(x):                                                # [S, heads, D]
        x1, x2 = x[..., : cfg.head_dim // 2], x[..., cfg.head_dim // 2:]
        c, s = cos[:, None], sin[:, None]
        return torch.cat([x1 * c - x2 * s, x2 * c + x1 * s], -1)

    h = hf["model.embed_tokens.weight"][ids[0]]
    for i in range(cfg.num_hidden_layers):
        p = f"model.layers.{i}."
        x = rms(h, hf[p + "input_layernorm.weight"])
        q = rope((x @ hf[p + "self_attn.q_proj.weight"].t()).view(10, cfg.num_attention_heads, cfg.head_dim))
        k = rope((x @ hf[p + "self_attn.k_proj.weight"].t()).view(10, cfg.num_key_value_heads, cfg.head_dim))
        v = (x @ hf[p + "self_attn.v_proj.weight"].t()).view(10, cfg.num_key_value_heads, cfg.head_dim)
        g = cfg.num_attention_heads // cfg.num_key_value_heads
        k, v = k.repeat_interleave(g, 1), v.repeat_interleave(g, 1)
        s = torch.einsum("shd,thd->hst", q, k) / cfg.head_dim ** 0.5
        s = s.masked_fill(~torch.ones(10, 10, dtype=torch.bool).tril(), float("-inf"))
        o = torch.einsum("hst,thd->shd", torch.softmax(s, -1), v).reshape(10, -1)
        h = h + o @ hf[p + "self_attn.o_proj.weight"].t()
        x = rms(h, hf[p + "post_attention_layernorm.weight"])
        h = h + (torch.nn.functional.silu(x @ hf[p + "mlp.gate_proj.weight"].t()) * (x @ hf[p + "mlp.up_proj.weight"].t())) \
            @ hf[p + "mlp.down_proj.weight"].t()
    ref = rms(h, hf["model.norm.weight"]) @ hf["model.embed_tokens.weight"].t()
    torch.testing.assert_close(logits, ref, atol=2e-4, rtol=1e-3)


def test_cli_paths_hooks_and_xser_conversion(tmp_path):
    """The reference-named steps: full -> sharded through the CLI entry point with a subclass hook, sharded torch.save ->
    xser -> torch.save, and merge back to the original full state."""
    import json

    hf = _hf_state(tied=False)
    src = tmp_path / "hf"
    os.makedirs(src)
    torch.save(hf, src / "pytorch_model.bin")
    cfg = tmp_path / "config.json"
    cfg.write_text(json.dumps({"num_hidden_layers": 4, "num_attention_heads": 8, "num_key_value_heads": 2,
                               "hidden_size": 64, "head_dim": 8, "tie_word_embeddings": False}))
    seen = {}

    class Conv(CheckpointConverterBase):
        def pre_process_full_state_before_tp_conversion(self, state_dict, args=None):
            seen["pre"] = len(state_dict)
            return state_dict

        def post_process_full_state_after_tp_conversion(self, state_dict, args=None):
            seen["post"] = len(state_dict)
            return state_dict

    common = ["--config", str(cfg), "--tp_size", "2", "--pp_size", "1"]
    Conv.run(["--input_dir", str(src), "--output_dir", str(tmp_path / "sharded"), "--convert_from_full_state"] + common)
    assert seen["pre"] == len(hf)
    Conv.run(["--input_dir", str(tmp_path / "sharded"), "--output_dir", str(tmp_path / "xser"), "--convert_to_xser"] + common)
    assert os.path.isdir(tmp_path / "xser" / "model" / "dp_rank_00_tp_rank_01_pp_rank_00.pt.tensors")
    Conv.run(["--input_dir", str(tmp_path / "xser"), "--output_dir", str(tmp_path / "plain"), "--convert_from_xser"] + common)
    Conv.run(["--input_dir", str(tmp_path / "plain"), "--output_dir", str(tmp_path / "full"), "--convert_to_full_state"]
             + common)
    assert seen["post"] == len(hf)
    back = torch.load(tmp_path / "full" / "checkpoint.pt", weights_only=False)
    assert set(back) == set(hf) and all(torch.equal(back[k], hf[k]) for k in hf)
    conv = Conv(4, 8, 2, 8, tp_size=2)
    assert conv.get_partition_dim("model.layers.0.mlp.down_proj.weight") == 1 and conv.get_partition_dim("model.norm.weight") is None
    assert conv.is_qkv_weight("model.layers.1.self_attn.k_proj.weight") and not conv.is_qkv_weight("model.norm.weight")
    fused = conv.coalesce_qkv(dict(hf))
    assert fused[conv.get_fused_qkv_key("model.layers.0.")].shape == (8 * 8 + 2 * 2 * 8, 64)
    shard = torch.load(tmp_path / "plain" / "model" / "dp_rank_00_tp_rank_00_pp_rank_00.pt", weights_only=False)
    split = conv.convert_partial_state_to_non_fused_qkv(dict(shard))
    assert split["model.layers.0.self_attn.qkv_proj.weight_q"].shape[0] == 32
    again = conv.convert_partial_state_to_fused_qkv(split)
    assert torch.equal(again["model.layers.0.self_attn.qkv_proj.weight_qkv"], shard["model.layers.0.self_attn.qkv_proj.weight_qkv"])
    assert conv.find_size({"a": torch.zeros(4, 4), "b": [torch.zeros(2, dtype=torch.bfloat16)]}) == 68


def test_megatron_style_round_trip(tmp_path):
    """``--model_style megatron``: a Megatron/NxDT-named full checkpoint (``query`` + ``key_value`` for GQA, fused
    ``dense_h_to_4h``) -> TP shards identical to those of the HF-named state -> back to Megatron names."""
    import json
    from types import SimpleNamespace

    hf = _hf_state(tied=False)
    conv = CheckpointConverterBase(4, 8, 2, 8, tp_size=2, tie_word_embeddings=False)
    meg = conv.hf_full_to_megatron(hf, qkv_linear=True)
    assert "language_model.embedding.word_embeddings.weight" in meg and "language_model.output_layer.weight" in meg
    assert "language_model.encoder.final_layernorm.weight" in meg
    p = "language_model.encoder.layers.1."
    assert meg[p + "self_attention.key_value.weight"].shape == (2 * 2 * 8, 64)
    assert meg[p + "mlp.dense_h_to_4h.weight"].shape == (192, 64) and p + "self_attention.dense.weight" in meg
    assert p + "input_layernorm.weight" in meg and not any(k.startswith("model.") for k in meg)
    back = conv.megatron_full_to_hf(meg)
    assert set(back) == set(hf) and all(torch.equal(back[k], hf[k]) for k in hf)
    mha = CheckpointConverterBase(4, 8, 8, 8)
    hf_mha = _hf_state(nkv=8, tied=False)
    meg_mha = mha.hf_full_to_megatron(hf_mha)
    assert meg_mha[p + "self_attention.query_key_value.weight"].shape == (3 * 64, 64)
    assert all(torch.equal(v, hf_mha[k]) for k, v in mha.megatron_full_to_hf(meg_mha).items())
    # single-key / in-place helpers under the reference's names
    assert conv.rename_keys_for_megatron("model.layers.0.mlp.down_proj.weight", "megatron", hf_to_nxdt=True) == \
        "language_model.encoder.layers.0.mlp.dense_4h_to_h.weight"
    assert conv.rename_keys_for_megatron("model.norm.weight", "hf", hf_to_nxdt=True) == "model.norm.weight"
    a = SimpleNamespace(model_style="megatron", qkv_linear=True, convert_from_full_state=False)
    part = {"x.self_attention.query.weight": torch.ones(4, 2), "x.self_attention.key_value.weight": torch.arange(8.).reshape(4, 2)}
    conv.modify_qkv_for_megatron(part, a)
    assert set(part) == {"x.self_attention.qkv_proj.weight_q", "x.self_attention.qkv_proj.weight_k",
                         "x.self_attention.qkv_proj.weight_v"}
    assert conv.is_q_or_o_for_megatron(a, "layers.0.self_attn.o_proj.weight")
    assert not conv.is_q_or_o_for_megatron(SimpleNamespace(model_style="hf"), "layers.0.self_attn.o_proj.weight")
    # CLI: Megatron full -> TP2 shards (same tensors as from HF names) -> Megatron full
    src = tmp_path / "meg"
    os.makedirs(src)
    torch.save(meg, src / "checkpoint.pt")
    cfg = tmp_path / "config.json"
    cfg.write_text(json.dumps({"num_hidden_layers": 4, "num_attention_heads": 8, "num_key_value_heads": 2,
                               "hidden_size": 64, "head_dim": 8, "tie_word_embeddings": False}))
    common = ["--config", str(cfg), "--tp_size", "2", "--model_style", "megatron"]
    CheckpointConverterBase.run(["--input_dir", str(src), "--output_dir", str(tmp_path / "sh"), "--convert_from_full_state"] + common)
    shard = torch.load(tmp_path / "sh" / "model" / "dp_rank_00_tp_rank_01_pp_rank_00.pt", weights_only=False)
    want = conv.shard_tp(conv.hf_to_nxd_full(hf), 1)
    assert set(shard) == set(want) and all(torch.equal(shard[k], want[k]) for k in want)
    CheckpointConverterBase.run(["--input_dir", str(tmp_path / "sh"), "--output_dir", str(tmp_path / "full"), "--convert_to_full_state"] + common)
    full = torch.load(tmp_path / "full" / "checkpoint.pt", weights_only=False)
    assert set(full) == set(meg) and all(torch.equal(full[k], meg[k]) for k in meg)


def test_mixtral_converter_tp_pp_ep_round_trip(tmp_path):
    """HF Mixtral names -> (tp 2, pp 2, ep 2) shards cut by the model's own parameter tags -> back, bit-exact; every
    shard has exactly the local shapes the parallel model allocates at that coordinate."""
    from neuronx_distributed_llama3_2_b200.models import (MixtralForCausalLM, convert_mixtral_state_dict_to_hf,
                                                          tiny_mixtral)
    from neuronx_distributed_llama3_2_b200.scripts.checkpoint_converter_moe import (CheckpointConverterMixtral,
                                                                                      mock_parallel_state)

    torch.manual_seed(0)
    cfg = tiny_mixtral(sequence_parallel_enabled=False, num_hidden_layers=4)
    full_model = MixtralForCausalLM(cfg)
    hf = convert_mixtral_state_dict_to_hf(full_model.state_dict(), cfg)
    conv = CheckpointConverterMixtral(cfg, tp_size=2, pp_size=2, ep_size=2)
    out = str(tmp_path / "sharded")
    conv.convert_from_full_state(hf, out)
    seen = {}
    for e in range(2):
        for t in range(2):
            with mock_parallel_state(2, t, 2, e), torch.device("meta"):
                local = MixtralForCausalLM(cfg, device="meta").state_dict()
            merged = {}
            for s in range(2):
                part = dict(torch.load(conv.shard_file(out, t, s, e), weights_only=False))      # expert tensors
                assert part and all("expert_mlps" in k for k in part)
                part.update(torch.load(conv.shard_file(out, t, s), weights_only=False))          # dense remainder
                assert all(("layers.0." in k or "layers.1." in k or "embed" in k) == (s == 0) for k in part
                           if "layers." in k or "embed" in k), (s, list(part))
                merged.update(part)
            assert set(merged) == set(local)
            for k, v in merged.items():
                assert tuple(v.shape) == tuple(local[k].shape), (k, v.shape, local[k].shape)
            seen[(e, t)] = merged
    w = full_model.state_dict()["layers.2.mlp.expert_mlps.mlp_op.down_proj.weight"]          # [E, I, H]: EP dim 0, TP dim 1
    I = w.shape[1]
    torch.testing.assert_close(seen[(1, 0)]["layers.2.mlp.expert_mlps.mlp_op.down_proj.weight"], w[2:, : I // 2])
    gu = full_model.state_dict()["layers.2.mlp.expert_mlps.mlp_op.gate_up_proj.weight"]      # [E, H, 2I], stride 2 on dim 2
    g, u = gu.chunk(2, dim=2)
    torch.testing.assert_close(seen[(0, 1)]["layers.2.mlp.expert_mlps.mlp_op.gate_up_proj.weight"],
                               torch.cat([g[:2, :, I // 2:], u[:2, :, I // 2:]], dim=2))
    back = conv.convert_to_full_state(out)
    assert set(back) == set(hf) and all(torch.equal(back[k], hf[k]) for k in hf)
    # CLI
    import json
    cfgp = tmp_path / "config.json"
    cfgp.write_text(json.dumps({"vocab_size": 512, "hidden_size": 128, "intermediate_size": 256, "num_hidden_layers": 4,
                                "num_attention_heads": 4, "num_key_value_heads": 2, "head_dim": 32, "num_local_experts": 4,
                                "num_experts_per_tok": 2, "tie_word_embeddings": False, "max_position_embeddings": 512}))
    torch.save(hf, tmp_path / "hf.pt")
    common = ["--config", str(cfgp), "--tp_size", "2", "--ep_size", "2"]
    CheckpointConverterMixtral.run(["--input_dir", str(tmp_path / "hf.pt"), "--output_dir", str(tmp_path / "s2"),
                                    "--convert_from_full_state", "--save_xser"] + common)
    CheckpointConverterMixtral.run(["--input_dir", str(tmp_path / "s2"), "--output_dir", str(tmp_path / "f2"),
                                    "--convert_to_full_state"] + common)
    again = torch.load(tmp_path / "f2" / "checkpoint.pt", weights_only=False)
    assert all(torch.equal(again[k], hf[k]) for k in hf)


def _load_converted_worker(rank, world, tmp):
    """HF-named weights -> offline converter (tp 2 x pp 2) -> ``nxd.load_checkpoint`` into a TP2 x PP2 (+SP) model: the
    training loss equals the one of a single-process model holding the same weights."""
    from neuronx_distributed_llama3_2_b200 import trainer as nxd
    from neuronx_distributed_llama3_2_b200.inference import _mock_parallel_state
    from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM, tiny_llama
    from neuronx_distributed_llama3_2_b200.models.llama import LlamaDecoderLayer

    ids = torch.randint(0, 512, (4, 32), generator=torch.Generator().manual_seed(9))
    with _mock_parallel_state(1, 0):
        torch.manual_seed(123)
        full = LlamaForCausalLM(tiny_llama(num_hidden_layers=4, sequence_parallel_enabled=False))
        conv = CheckpointConverterBase(4, 8, 2, 32, tp_size=2, pp_size=2, tie_word_embeddings=True)
        hf = CheckpointConverterBase(4, 8, 2, 32, tie_word_embeddings=True).nxd_full_to_hf(full.state_dict())
        with torch.no_grad():
            want = float(torch.stack([full(c, labels=c) for c in torch.chunk(ids, 2, 0)]).mean())
    ckpt = os.path.join(tmp, "pretrained")
    if rank == 0:
        conv.convert_from_full_state(hf, ckpt)
    torch.distributed.barrier()
    nxd_cfg = nxd.neuronx_distributed_config(
        tensor_parallel_size=2, pipeline_parallel_size=2, sequence_parallel=True,
        pipeline_config={"transformer_layer_cls": LlamaDecoderLayer, "num_microbatches": 2, "auto_partition": True,
                         "input_names": ["input_ids", "labels"], "broadcast_and_average_loss": True})
    torch.manual_seed(7)                                  # different random init: the checkpoint must overwrite it
    model = nxd.initialize_parallel_model(nxd_cfg, lambda: LlamaForCausalLM(tiny_llama(num_hidden_layers=4)),
                                          device=torch.device("cpu"))
    nxd.load_checkpoint(tmp, tag="pretrained", model=model, strict=False)
    got = float(model.run_train(input_ids=ids, labels=ids))
    assert abs(got - want) < 2e-5, (rank, got, want)


def test_converted_tp2_pp2_checkpoint_loads_into_parallel_model(tmp_path):
    from conftest import run_distributed

    run_distributed(_load_converted_worker, 4, str(tmp_path))


def _load_kv_replicated_worker(rank, world, tmp):
    """tp = 4 with 2 KV heads: the converter tiles the KV heads twice and re-orders the query heads so that every rank's
    queries sit next to the KV head it holds; the model (``kv_shared_group_size=2``) must compute the same loss."""
    from neuronx_distributed_llama3_2_b200 import trainer as nxd
    from neuronx_distributed_llama3_2_b200.inference import _mock_parallel_state
    from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM, tiny_llama

    ids = torch.randint(0, 512, (2, 32), generator=torch.Generator().manual_seed(9))
    with _mock_parallel_state(1, 0):
        torch.manual_seed(321)
        full = LlamaForCausalLM(tiny_llama(sequence_parallel_enabled=False))
        hf = CheckpointConverterBase(2, 8, 2, 32, tie_word_embeddings=True).nxd_full_to_hf(full.state_dict())
        with torch.no_grad():
            want = float(full(ids, labels=ids))
    ckpt = os.path.join(tmp, "pretrained")
    if rank == 0:
        CheckpointConverterBase(2, 8, 2, 32, tp_size=4, kv_size_multiplier=2, tie_word_embeddings=True
                                ).convert_from_full_state(hf, ckpt)
    torch.distributed.barrier()
    nxd_cfg = nxd.neuronx_distributed_config(tensor_parallel_size=4, sequence_parallel=True)
    torch.manual_seed(7)
    model = nxd.initialize_parallel_model(nxd_cfg, lambda: LlamaForCausalLM(tiny_llama(kv_shared_group_size=2)),
                                          device=torch.device("cpu"))
    nxd.load_checkpoint(tmp, tag="pretrained", model=model, strict=False)
    got = float(model(ids, labels=ids))
    assert abs(got - want) < 2e-5, (rank, got, want)


def test_converted_kv_replicated_checkpoint_loads_at_tp4(tmp_path):
    from conftest import run_distributed

    run_distributed(_load_kv_replicated_worker, 4, str(tmp_path))
