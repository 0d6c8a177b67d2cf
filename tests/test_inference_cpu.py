"""Inference path on CPU: KV-cache generation equals full recomputation, ModelBuilder shape routing, per-rank sharded
safetensors (reference ``test/integration/inference/test_model_builder.py`` + runner accuracy check)."""
from __future__ import annotations

import os
import tempfile

import pytest

import torch

from neuronx_distributed_llama3_2_b200.inference import (BaseModelInstance, InferenceConfig, LlamaInferenceModel,
                                                         ModelBuilder, Sampler, get_sharded_checkpoint,
                                                         _mock_parallel_state)
from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM, tiny_llama


def _models(paged=False):
    torch.manual_seed(0)
    cfg = tiny_llama(sequence_parallel_enabled=False)
    train = LlamaForCausalLM(cfg)
    inf = LlamaInferenceModel(tiny_llama(sequence_parallel_enabled=False),
                              InferenceConfig(batch_size=2, max_context_length=32, seq_len=64, paged_kv=paged,
                                              page_size=16))
    inf.load_state_dict(train.state_dict())
    return cfg, train, inf


def _full_logits(train, ids):
    """[B, T, V] logits of the training model (sequence-first inside)."""
    with torch.no_grad():
        return train(ids).transpose(0, 1)


def test_kv_cache_generation_matches_full_recompute():
    for paged in (False, True):
        cfg, train, inf = _models(paged)
        prompts = torch.randint(1, cfg.vocab_size, (2, 12))
        mask = torch.ones_like(prompts)
        mask[1, 9:] = 0                      # second prompt is shorter (right padding)
        out = inf.generate(prompts, mask, max_new_tokens=6)
        # oracle: greedy decoding by recomputing the whole sequence each step with the training model
        for b in range(2):
            L = int(mask[b].sum())
            seq = prompts[b, :L].clone()
            for _ in range(6):
                nxt = _full_logits(train, seq.unsqueeze(0))[0, -1].argmax()
                seq = torch.cat([seq, nxt.reshape(1)])
            assert torch.equal(out[b, :L + 6], seq), (paged, b, out[b, :L + 6], seq)


def test_prefill_logits_match_training_model():
    cfg, train, inf = _models()
    inf.init_kv_cache()
    ids = torch.randint(1, cfg.vocab_size, (2, 16))
    pos = torch.arange(16).unsqueeze(0).expand(2, 16)
    tok, logits = inf.forward_infer(ids, pos, return_logits=True)
    ref = _full_logits(train, ids)[:, -1]
    torch.testing.assert_close(logits, ref, atol=1e-4, rtol=1e-4)
    assert torch.equal(tok, ref.argmax(-1))


def test_sampler_topk_reproducible():
    s = Sampler(top_k=5, do_sample=True, temperature=0.7, seed=3)
    logits = torch.randn(4, 100)
    u = torch.rand(4)
    a, b = s.sample(logits, u), s.sample(logits, u)
    assert torch.equal(a, b)
    top5 = logits.topk(5, -1).indices
    assert all(int(a[i]) in top5[i].tolist() for i in range(4))
    assert torch.equal(Sampler(top_k=1).sample(logits), logits.argmax(-1))


def test_model_builder_routes_by_shape_and_shards_checkpoint():
    cfg, train, _ = _models()
    full_sd = {k: v.clone() for k, v in train.state_dict().items()}

    def make():
        m = LlamaInferenceModel(tiny_llama(sequence_parallel_enabled=False),
                                InferenceConfig(batch_size=2, max_context_length=32, seq_len=64))
        m.init_kv_cache(device=torch.device("cpu"))
        return m

    builder = ModelBuilder(tp_degree=1, checkpoint_loader=lambda: full_sd, use_cuda_graphs=False)
    cte = [(torch.zeros(2, 16, dtype=torch.long), torch.zeros(2, 16, dtype=torch.long)),
           (torch.zeros(2, 32, dtype=torch.long), torch.zeros(2, 32, dtype=torch.long))]
    shared = BaseModelInstance(make)
    builder.add("context_encoding_model", shared, cte, priority_model_idx=0)
    builder.add("token_generation_model", shared, [(torch.zeros(2, 1, dtype=torch.long), torch.zeros(2, 1, dtype=torch.long))])
    nxd_model = builder.trace()
    assert set(nxd_model.keys()) == {"context_encoding_model", "token_generation_model"}
    ids = torch.randint(1, cfg.vocab_size, (2, 16))
    pos = torch.arange(16).unsqueeze(0).expand(2, 16)
    tok = nxd_model(ids, pos)
    assert torch.equal(tok, _full_logits(train, ids)[:, -1].argmax(-1))
    tok2 = nxd_model(tok.reshape(2, 1), torch.full((2, 1), 16))
    ref2 = _full_logits(train, torch.cat([ids, tok.reshape(2, 1)], 1))[:, -1].argmax(-1)
    assert torch.equal(tok2, ref2)
    try:
        nxd_model(torch.zeros(3, 7, dtype=torch.long), torch.zeros(3, 7, dtype=torch.long))
        raise AssertionError("unrouted shape must fail")
    except RuntimeError:
        pass

    # per-rank sharded safetensors for tp=2 written from ONE process (mocked parallel state + meta modules)
    from safetensors.torch import load_file

    b2 = ModelBuilder(tp_degree=2, checkpoint_loader=lambda: full_sd)
    b2.add("m", BaseModelInstance(lambda: LlamaForCausalLM(tiny_llama(sequence_parallel_enabled=False))), [])
    with tempfile.TemporaryDirectory() as d:
        b2.shard_checkpoint(d)
        shards = [load_file(os.path.join(d, f"tp{r}_sharded_checkpoint.safetensors")) for r in range(2)]
    w = full_sd["model.layers.0.mlp.down_proj.weight"]             # row parallel: split on dim 1
    torch.testing.assert_close(torch.cat([s["model.layers.0.mlp.down_proj.weight"] for s in shards], 1), w)
    gu = full_sd["model.layers.0.mlp.gate_up_proj.weight"]         # stride-2 column parallel: [g0,u0 | g1,u1]
    g, u = gu.chunk(2, 0)
    torch.testing.assert_close(shards[1]["model.layers.0.mlp.gate_up_proj.weight"],
                               torch.cat([g.chunk(2, 0)[1], u.chunk(2, 0)[1]], 0))
    qkv = full_sd["model.layers.0.self_attn.qkv_proj.weight_qkv"]  # fused QKV: each part sharded on its own
    q, k, v = torch.split(qkv, [8 * 32, 2 * 32, 2 * 32], 0)
    torch.testing.assert_close(shards[0]["model.layers.0.self_attn.qkv_proj.weight_qkv"],
                               torch.cat([q.chunk(2, 0)[0], k.chunk(2, 0)[0], v.chunk(2, 0)[0]], 0))
    torch.testing.assert_close(shards[0]["model.norm.weight"], full_sd["model.norm.weight"])
    # the same slicing through the explicit API
    with _mock_parallel_state(2, 1):
        with torch.device("meta"):
            meta = LlamaForCausalLM(tiny_llama(sequence_parallel_enabled=False))
        s1 = get_sharded_checkpoint(full_sd, meta, 1, 2)
    torch.testing.assert_close(s1["model.embed_tokens.weight"], full_sd["model.embed_tokens.weight"].chunk(2, 0)[1])


def test_runner_generate_buckets_and_benchmark_report():
    from neuronx_distributed_llama3_2_b200.inference import InferenceRunner

    cfg, train, _ = _models()
    sd = {k: v.clone() for k, v in train.state_dict().items()}
    ic = InferenceConfig(batch_size=2, max_context_length=64, seq_len=128, buckets=[16, 64], max_new_tokens=5,
                         use_cuda_graphs=False)
    runner = InferenceRunner(tiny_llama(sequence_parallel_enabled=False), ic, checkpoint_loader=lambda: sd,
                             device=torch.device("cpu"))
    with tempfile.TemporaryDirectory() as d:
        runner.trace(d)
        assert os.path.exists(os.path.join(d, "weights", "tp0_sharded_checkpoint.safetensors"))
        assert os.path.exists(os.path.join(d, "config.json"))
    prompts = torch.randint(1, cfg.vocab_size, (2, 20))       # length 20 -> padded to the 64 bucket
    mask = torch.ones_like(prompts)
    mask[0, 13:] = 0
    out = runner.generate(prompts, mask, max_new_tokens=5)
    for b in range(2):
        L = int(mask[b].sum())
        seq = prompts[b, :L].clone()
        for _ in range(5):
            seq = torch.cat([seq, _full_logits(train, seq.unsqueeze(0))[0, -1].argmax().reshape(1)])
        assert torch.equal(out[b, :L + 5], seq)
    assert runner.check_accuracy_logits(train, prompts[:, :16], atol=1e-3)
    rep = runner.benchmark_sampling(prompt_len=16, new_tokens=4, n_runs=2)
    assert rep["ttft_ms_p50"] > 0 and rep["decode_tokens_per_s"] > 0
    assert set(rep["e2e_model"]) >= {"latency_ms_p50", "latency_ms_p99", "throughput"}
    with tempfile.TemporaryDirectory() as d:
        import json
        path = runner.torch_profile(os.path.join(d, "prof", "trace.json"), prompt_len=16, new_tokens=2)
        with open(path) as f:
            assert len(json.load(f)["traceEvents"]) > 0


def _attention(x, q, k, v, o, hq, hkv, d):
    """Plain GQA attention (no mask) from projection weights: the oracle for the head-rewrite check."""
    B, T, _ = x.shape
    Q = (x @ q.t()).reshape(B, T, hq, d).transpose(1, 2)
    K = (x @ k.t()).reshape(B, T, hkv, d).transpose(1, 2)
    V = (x @ v.t()).reshape(B, T, hkv, d).transpose(1, 2)
    rep = hq // hkv
    K, V = K.repeat_interleave(rep, 1), V.repeat_interleave(rep, 1)
    att = torch.softmax(Q @ K.transpose(-1, -2) / d ** 0.5, -1) @ V
    return att.transpose(1, 2).reshape(B, T, hq * d) @ o.t()


def test_gqa_sharding_strategies_preserve_attention():
    from neuronx_distributed_llama3_2_b200.inference.gqa import (GQA, determine_sharding_strategy,
                                                                  get_shardable_head_counts, transform_attention_weights)

    assert determine_sharding_strategy(32, 8) is GQA.REPLICATE_TO_TP_DEGREE
    assert determine_sharding_strategy(12, 8) is GQA.CONVERT_TO_MHA
    assert get_shardable_head_counts(32, 56, 8, GQA.REPLICATE_TO_TP_DEGREE) == (64, 32)
    assert get_shardable_head_counts(32, 56, 8, GQA.CONVERT_TO_MHA) == (64, 64)
    assert get_shardable_head_counts(8, 32, 8, GQA.REPLICATE_TO_TP_DEGREE) == (32, 8)
    torch.manual_seed(0)
    hidden, d = 48, 8
    x = torch.randn(2, 5, hidden, dtype=torch.float64)
    for hq, hkv, tp, want in ((6, 2, 4, GQA.REPLICATE_TO_TP_DEGREE), (6, 2, 3, GQA.CONVERT_TO_MHA),
                              (6, 6, 4, GQA.REPLICATE_TO_TP_DEGREE), (8, 2, 2, GQA.REPLICATE_TO_TP_DEGREE)):
        q = torch.randn(hq * d, hidden, dtype=torch.float64)
        k = torch.randn(hkv * d, hidden, dtype=torch.float64)
        v = torch.randn(hkv * d, hidden, dtype=torch.float64)
        o = torch.randn(hidden, hq * d, dtype=torch.float64)
        ref = _attention(x, q, k, v, o, hq, hkv, d)
        r = transform_attention_weights(q, k, v, o, tp, hq, hkv)
        nq, nkv = r["num_attention_heads"], r["num_key_value_heads"]
        assert nq % tp == 0 and (nkv % tp == 0) and nq % nkv == 0, (hq, hkv, tp, nq, nkv)
        if hq != hkv and tp % hkv == 0 and hkv < tp:
            assert r["strategy"] is want and nkv == tp
        got = _attention(x, r["q"], r["k"], r["v"], r["o"], nq, nkv, d)
        torch.testing.assert_close(got, ref, atol=1e-9, rtol=1e-9)


def test_autobucketing_generate_and_fit():
    from neuronx_distributed_llama3_2_b200.inference.autobucketing import (context_encoder_bk, generate_buckets,
                                                                            select_bucket, token_generation_bk)

    assert generate_buckets(128, 2048) == [128, 256, 512, 1024, 2048]
    assert generate_buckets(128, 1100) == [128, 256, 512, 1100]
    assert generate_buckets(128, 1500) == [128, 256, 512, 1024, 1500]
    assert generate_buckets(64, 64) == [64]
    buckets = [16, 32, 64]
    assert select_bucket(buckets, 16) == 0 and select_bucket(buckets, 17) == 1
    ids = torch.arange(1, 21).reshape(1, 20).repeat(2, 1)
    mask = torch.ones(2, 20, dtype=torch.long)
    mask[1, 12:] = 0
    pos = torch.arange(20).unsqueeze(0).repeat(2, 1)
    (i2, m2, p2), idx = context_encoder_bk([ids, mask, pos], buckets, "right", pad_token=0)
    assert idx == 1 and i2.shape == (2, 32) and torch.equal(i2[:, :20], ids) and int(i2[:, 20:].abs().sum()) == 0
    assert int(m2[:, 20:].sum()) == 0 and torch.equal(p2[0], torch.arange(32))
    (i3, _, _), idx = context_encoder_bk([ids[:, :20], torch.cat([mask[:, :10], mask[:, :10] * 0], 1), pos], buckets)
    assert idx == 0 and i3.shape == (2, 16)                       # longest real prompt is 10 tokens: sliced to bucket 16
    cache_mask = torch.ones(2, 64, dtype=torch.long)
    (_, m4, _), idx = token_generation_bk([ids[:, :1], cache_mask, torch.tensor([[20], [12]])], buckets)
    assert idx == 1 and m4.shape == (2, 32)


def test_hf_checkpoint_directory_formats(tmp_path):
    from neuronx_distributed_llama3_2_b200.inference.checkpoint import load_state_dict, save_state_dict_safetensors

    torch.manual_seed(0)
    sd = {f"layer.{i}.w": torch.randn(64, 32) for i in range(6)}
    sd["tied"] = sd["layer.0.w"]
    one = tmp_path / "one"
    save_state_dict_safetensors(sd, str(one))
    assert sorted(os.listdir(one)) == ["model.safetensors"]
    back = load_state_dict(str(one))
    assert "tied" not in back and all(torch.equal(back[k], sd[k]) for k in back) and len(back) == 6
    many = tmp_path / "many"
    save_state_dict_safetensors(sd, str(many), max_shard_size="20KB")          # 8 KB per tensor -> 2 per shard
    names = sorted(os.listdir(many))
    assert "model.safetensors.index.json" in names and sum(n.endswith(".safetensors") for n in names) == 3
    back = load_state_dict(str(many))
    assert len(back) == 6 and all(torch.equal(back[k], sd[k]) for k in back)
    binv = tmp_path / "bin"
    os.makedirs(binv)
    torch.save({k: v for k, v in sd.items() if k != "tied"}, binv / "pytorch_model.bin")
    assert torch.equal(load_state_dict(str(binv))["layer.3.w"], sd["layer.3.w"])
    try:
        load_state_dict(str(tmp_path))
        raise AssertionError("empty directory must fail")
    except FileNotFoundError:
        pass


def test_mixtral_kv_cache_generation_matches_full_recompute():
    from neuronx_distributed_llama3_2_b200.inference import MixtralInferenceModel
    from neuronx_distributed_llama3_2_b200.models import MixtralForCausalLM, tiny_mixtral

    torch.manual_seed(0)
    cfg = tiny_mixtral(sequence_parallel_enabled=False)
    train = MixtralForCausalLM(cfg).eval()
    inf = MixtralInferenceModel(tiny_mixtral(sequence_parallel_enabled=False),
                                InferenceConfig(batch_size=2, max_context_length=32, seq_len=64))
    inf.load_state_dict(train.state_dict())
    prompts = torch.randint(1, cfg.vocab_size, (2, 10))
    mask = torch.ones_like(prompts)
    mask[1, 7:] = 0
    out = inf.generate(prompts, mask, max_new_tokens=5)
    for b in range(2):
        L = int(mask[b].sum())
        seq = prompts[b, :L].clone()
        for _ in range(5):
            with torch.no_grad():
                nxt = train(seq.unsqueeze(0)).transpose(0, 1)[0, -1].argmax()
            seq = torch.cat([seq, nxt.reshape(1)])
        assert torch.equal(out[b, :L + 5], seq), (b, out[b, :L + 5], seq)


def test_model_wrapper_pads_splits_and_reorders():
    from neuronx_distributed_llama3_2_b200.inference.model_wrapper import ModelWrapper

    cfg, train, _ = _models()
    ic = InferenceConfig(batch_size=2, max_batch_size=5, max_context_length=32, seq_len=64, is_continuous_batching=True)
    inf = LlamaInferenceModel(tiny_llama(sequence_parallel_enabled=False), ic)
    inf.load_state_dict(train.state_dict())
    inf.init_kv_cache()
    calls = []

    def fn(ids, pos, seq, last):
        calls.append((tuple(ids.shape), seq.tolist()))
        return inf.forward_infer(ids, pos, seq, last)

    cte = ModelWrapper(fn, batch_size=2, buckets=[16, 32], is_continuous_batching=True)
    tkg = ModelWrapper(fn, batch_size=2, buckets=[1], tag="token_generation_model", is_continuous_batching=True)
    prompts = torch.randint(1, cfg.vocab_size, (3, 11))              # 3 requests on a batch-2 graph, 11 tokens -> bucket 16
    pos = torch.arange(11).unsqueeze(0).expand(3, 11)
    lines = torch.tensor([4, 0, 2], dtype=torch.int32)               # cache lines out of order
    tok = cte(prompts, pos, lines)
    assert [c[0] for c in calls] == [(2, 16), (2, 16)]               # split 2 + 1 (the single row padded by repetition)
    assert calls[0][1] == [0, 2] and calls[1][1] == [4, 4]            # sorted by cache line
    ref = _full_logits(train, prompts)[:, -1].argmax(-1)
    assert torch.equal(tok, ref)                                      # caller's row order restored
    tok2 = tkg(tok.reshape(3, 1), torch.full((3, 1), 11), lines)
    ref2 = _full_logits(train, torch.cat([prompts, tok.reshape(3, 1)], 1))[:, -1].argmax(-1)
    assert torch.equal(tok2, ref2)
    try:
        cte(torch.zeros(1, 40, dtype=torch.long), torch.arange(40).unsqueeze(0))
        raise AssertionError("longer than the largest bucket must fail")
    except ValueError:
        pass


def test_model_builder_named_steps():
    cfg, train, _ = _models()
    full_sd = {k: v.clone() for k, v in train.state_dict().items()}

    def make():
        m = LlamaInferenceModel(tiny_llama(sequence_parallel_enabled=False),
                                InferenceConfig(batch_size=2, max_context_length=32, seq_len=64))
        m.init_kv_cache(device=torch.device("cpu"))
        return m

    builder = ModelBuilder(tp_degree=1, checkpoint_loader=lambda: full_sd, use_cuda_graphs=False)
    inst = BaseModelInstance(make)
    ex = (torch.zeros(2, 16, dtype=torch.long), torch.zeros(2, 16, dtype=torch.long))
    builder.add("context_encoding_model", inst, [ex], priority_model_idx=0)
    shard = builder.shard_weights(0, builder.model_collection["context_encoding_model"])
    assert set(shard) >= {"model.norm.weight", "lm_head.weight"}
    inst.load_module()
    inst.module.load_state_dict(shard, strict=False)
    model = builder.build_nxd_model({"context_encoding_model": inst.module})
    init = builder.build_state_initializer({"context_encoding_model": inst.module})
    state = init()
    assert any(k.endswith("kv_cache.k") for k in state) and all(v.abs().sum() == 0 for v in state.values())
    assert builder.build_flattener_map()["context_encoding_model"]([ex[0], (ex[1],)]) == [ex[0], ex[1]]
    assert builder.build_packer()([ex[0]]) is ex[0]
    ids = torch.randint(1, cfg.vocab_size, (2, 16))
    pos = torch.arange(16).unsqueeze(0).expand(2, 16)
    assert torch.equal(model(ids, pos), _full_logits(train, ids)[:, -1].argmax(-1))
    model.mock_initialization(True)
    assert model.weights_initialized
    with tempfile.TemporaryDirectory() as d:
        builder.shard_checkpoint(d)
        with torch.no_grad():
            inst.module.model.norm.weight.zero_()
        model.initialize_with_saved_weights(d, rank=0)
    assert torch.equal(model(ids, pos), _full_logits(train, ids)[:, -1].argmax(-1))
    # the bucket-kernel front end of the reference's SPMD wrappers: pad to a bucket, pick its index, run it
    from neuronx_distributed_llama3_2_b200.inference.model_builder import (SPMDBucketModel, SPMDBucketModelScript,
                                                                            _Bucket)

    script = SPMDBucketModelScript([_Bucket(inst.module, (ids[:, :8], pos[:, :8]), False),
                                    _Bucket(inst.module, (ids, pos), False)])

    def kernel(inputs, buckets):
        T = inputs[0].shape[1]
        idx = next(i for i, b in enumerate(buckets) if T <= b)
        return inputs, torch.tensor(idx)

    spmd = SPMDBucketModel(kernel, ([8, 16],), script)
    assert torch.equal(spmd([ids, pos]), model(ids, pos))
    with pytest.raises(ValueError):
        script([ids, pos], torch.tensor(5))


def test_dbrx_generation_matches_recompute_and_hf_key_conversion():
    from neuronx_distributed_llama3_2_b200.inference import DbrxInferenceModel
    from neuronx_distributed_llama3_2_b200.models import DbrxForCausalLM, convert_dbrx_hf_state_dict, tiny_dbrx

    torch.manual_seed(0)
    cfg = tiny_dbrx(sequence_parallel_enabled=False)
    train = DbrxForCausalLM(cfg).eval()
    with torch.no_grad():                                  # make the clamp bite: some projections exceed clip_qkv = 0.5
        for layer in train.layers:
            layer.self_attn.qkv_proj.weight_qkv.mul_(8.0)
    inf = DbrxInferenceModel(tiny_dbrx(sequence_parallel_enabled=False),
                             InferenceConfig(batch_size=2, max_context_length=32, seq_len=64))
    inf.load_state_dict(train.state_dict())
    prompts = torch.randint(1, cfg.vocab_size, (2, 9))
    out = inf.generate(prompts, max_new_tokens=4)
    for b in range(2):
        seq = prompts[b].clone()
        for _ in range(4):
            with torch.no_grad():
                nxt = train(seq.unsqueeze(0)).transpose(0, 1)[0, -1].argmax()
            seq = torch.cat([seq, nxt.reshape(1)])
        assert torch.equal(out[b], seq), (b, out[b], seq)
    # HF DBRX key layout -> ours (experts stacked [E * I, H] there, [E, H, 2I] / [E, I, H] here)
    E, I, H = cfg.num_local_experts, cfg.intermediate_size, cfg.hidden_size
    sd = train.state_dict()
    hf = {"transformer.wte.weight": sd["embed_tokens.weight"], "transformer.norm_f.weight": sd["norm.weight"],
          "lm_head.weight": sd["lm_head.weight"]}
    for i in range(cfg.num_hidden_layers):
        p, q = f"layers.{i}.", f"transformer.blocks.{i}."
        gu = sd[p + "mlp.expert_mlps.mlp_op.gate_up_proj.weight"]                       # [E, H, 2I]
        hf[q + "ffn.experts.mlp.w1"] = gu[:, :, :I].transpose(1, 2).reshape(E * I, H)
        hf[q + "ffn.experts.mlp.v1"] = gu[:, :, I:].transpose(1, 2).reshape(E * I, H)
        hf[q + "ffn.experts.mlp.w2"] = sd[p + "mlp.expert_mlps.mlp_op.down_proj.weight"].reshape(E * I, H)
        hf[q + "ffn.router.layer.weight"] = sd[p + "mlp.router.linear_router.weight"]
        hf[q + "norm_attn_norm.attn.Wqkv.weight"] = sd[p + "self_attn.qkv_proj.weight_qkv"]
        hf[q + "norm_attn_norm.attn.out_proj.weight"] = sd[p + "self_attn.o_proj.weight"]
        hf[q + "norm_attn_norm.norm_1.weight"] = sd[p + "input_layernorm.weight"]
        hf[q + "norm_attn_norm.norm_2.weight"] = sd[p + "post_attention_layernorm.weight"]
    back = convert_dbrx_hf_state_dict(hf, cfg)
    assert set(back) == set(sd) and all(torch.equal(back[k], sd[k]) for k in sd)


def test_family_runners_and_quantized_runner():
    """``MixtralRunner`` drives the MoE KV-cache model through the same trace / generate path; a quantized Llama runner
    swaps the decoder linears for int8 weight-only layers, quantizes the float checkpoint on load and stays close to
    the float model."""
    from neuronx_distributed_llama3_2_b200.inference import InferenceRunner, MixtralRunner
    from neuronx_distributed_llama3_2_b200.models import MixtralForCausalLM, tiny_mixtral
    from neuronx_distributed_llama3_2_b200.quantization.quantization_layers import BaseQuantizeParallelLinear

    torch.manual_seed(0)
    mcfg = tiny_mixtral(sequence_parallel_enabled=False)
    mtrain = MixtralForCausalLM(mcfg).eval()
    msd = {k: v.clone() for k, v in mtrain.state_dict().items()}
    ic = InferenceConfig(batch_size=2, max_context_length=32, seq_len=64, buckets=[16, 32], use_cuda_graphs=False)
    r = MixtralRunner(tiny_mixtral(sequence_parallel_enabled=False), ic, checkpoint_loader=lambda: msd,
                      device=torch.device("cpu"))
    r.trace()
    prompts = torch.randint(1, mcfg.vocab_size, (2, 9))
    out = r.generate(prompts, max_new_tokens=4)
    seq = prompts[0].clone()
    for _ in range(4):
        with torch.no_grad():
            seq = torch.cat([seq, mtrain(seq.unsqueeze(0)).transpose(0, 1)[0, -1].argmax().reshape(1)])
    assert torch.equal(out[0], seq)

    cfg, train, _ = _models()
    sd = {k: v.clone() for k, v in train.state_dict().items()}
    for qtype in ("per_tensor_symmetric", "per_channel_symmetric"):
        icq = InferenceConfig(batch_size=2, max_context_length=32, seq_len=64, buckets=[16, 32], use_cuda_graphs=False,
                              quantized=True, quantization_type=qtype)
        rq = InferenceRunner(tiny_llama(sequence_parallel_enabled=False), icq, checkpoint_loader=lambda: sd,
                             device=torch.device("cpu"))
        qsd = rq.generate_quantized_state_dict(sd)
        k = "model.layers.0.mlp.down_proj"
        assert qsd[k + ".weight"].dtype == torch.int8 and k + ".scale" in qsd
        assert qsd["model.layers.0.self_attn.qkv_proj.weight_qkv"].is_floating_point() and "lm_head.scale" not in qsd
        with tempfile.TemporaryDirectory() as d:
            rq.trace(d)
            assert os.path.exists(os.path.join(d, "weights", "tp0_sharded_checkpoint.safetensors"))
        n_q = sum(isinstance(m, BaseQuantizeParallelLinear) for m in rq.model.modules())
        assert n_q == 3 * cfg.num_hidden_layers, n_q                    # o_proj, gate_up_proj, down_proj per layer
        ids = torch.randint(1, cfg.vocab_size, (2, 16))
        pos = torch.arange(16).unsqueeze(0).expand(2, 16)
        _, logits = rq.model.forward_infer(ids, pos, return_logits=True)
        ref = _full_logits(train, ids)[:, -1]
        rel = ((logits - ref).norm() / ref.norm()).item()
        assert rel < 0.05, (qtype, rel)


def test_continuous_batching_over_paged_pool_matches_isolated_generation():
    """Six requests of different lengths through three cache lines and a page pool that cannot hold them all at once:
    admission waits for pages, finished requests hand theirs back, a starved request is preempted and re-run -- and every
    request still gets exactly the tokens it gets when generated alone on a contiguous cache."""
    from neuronx_distributed_llama3_2_b200.inference.continuous_batching import ContinuousBatcher
    from neuronx_distributed_llama3_2_b200.runtime.native import NativeScheduler, PagedKVPool

    # the allocator itself: growth, fork (copy-on-write of the partial page), un-share, truncate, exhaustion
    pool = PagedKVPool(6, 4, 3, 4)
    assert pool.reserve(0, 5) == 2 and pool.num_free == 4 and pool.block_row(0)[0][:2] == [0, 1]
    assert pool.fork(0, 1) == (1, 2) and pool.block_row(1)[0][:2] == [0, 2] and pool.page_shared(0, 0)
    assert not pool.page_shared(0, 4) and pool.unshare(1, 0) == (0, 3) and pool.unshare(1, 0) is None
    assert pool.truncate(0, 4) == 1 and pool.free_line(1) == 2 and pool.free_line(0) == 1 and pool.num_free == 6
    with pytest.raises(MemoryError):
        pool.reserve(2, 17)
    sched = NativeScheduler(pool, 2)
    for i, (pl, mn) in enumerate([(5, 3), (3, 2), (2, 2)]):
        sched.submit(100 + i, pl, mn)
    assert sched.admit() == [(100, 0), (101, 1)] and sched.num_waiting == 1
    assert sched.step() == ([], [], 2) and sched.step() == ([101], [], 1)
    assert sched.admit() == [(102, 1)] and sched.step() == ([100], [], 1) and sched.step() == ([102], [], 0)
    assert pool.num_free == 6

    cfg, train, _ = _models()
    sd = train.state_dict()
    ref_model = LlamaInferenceModel(tiny_llama(sequence_parallel_enabled=False),
                                    InferenceConfig(batch_size=1, max_context_length=32, seq_len=64))
    ref_model.load_state_dict(sd)
    g = torch.Generator().manual_seed(7)
    specs = [(11, 9), (5, 14), (17, 6), (8, 12), (3, 10), (13, 7)]
    prompts = [torch.randint(1, cfg.vocab_size, (T,), generator=g) for T, _ in specs]
    want = [ref_model.generate(p.unsqueeze(0), max_new_tokens=n)[0, p.numel():].tolist() for p, (_, n) in zip(prompts, specs)]
    for n_pages, expect_preempt in ((12, False), (4, True)):
        ic = InferenceConfig(batch_size=3, max_context_length=32, seq_len=64, paged_kv=True, page_size=8,
                             num_kv_pages=n_pages, is_continuous_batching=True)
        model = LlamaInferenceModel(tiny_llama(sequence_parallel_enabled=False), ic)
        model.load_state_dict(sd)
        engine = ContinuousBatcher(model, max_running=3, pad_decode_to=3 if n_pages == 12 else None)
        rids = [engine.submit(p, n) for p, (_, n) in zip(prompts, specs)]
        got = engine.run()
        for rid, w in zip(rids, want):
            assert got[rid] == w, (n_pages, rid, got[rid], w)
        st = engine.stats
        assert st["prefills"] == 6 + st["preemptions"] and st["max_running"] <= 3
        assert model.kv_cache.pool.num_free == n_pages
        assert st["decode_rows"] > st["decode_calls"]                       # requests really shared decode calls
        assert (st["preemptions"] > 0) == expect_preempt, st


def test_runner_with_dynamic_page_pool_reserves_lines():
    """``num_kv_pages`` with the runner: ``generate`` gives every cache line its pages up front (the loop itself never talks
    to the allocator) and successive calls recycle them."""
    from neuronx_distributed_llama3_2_b200.inference import InferenceRunner

    cfg, train, _ = _models()
    sd = {k: v.clone() for k, v in train.state_dict().items()}
    outs = []
    for kw in (dict(), dict(paged_kv=True, page_size=8, num_kv_pages=12)):
        ic = InferenceConfig(batch_size=2, max_context_length=32, seq_len=64, buckets=[16, 32], use_cuda_graphs=False, **kw)
        r = InferenceRunner(tiny_llama(sequence_parallel_enabled=False), ic, checkpoint_loader=lambda: sd, device=torch.device("cpu"))
        r.trace()
        p = torch.randint(1, cfg.vocab_size, (2, 11), generator=torch.Generator().manual_seed(2))
        outs.append([r.generate(p, max_new_tokens=9), r.generate(p[:, :7], max_new_tokens=20)])
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    assert r.model.kv_cache.pool.num_free == 12 - 2 * ((16 + 20 + 7) // 8)
