"""The inference stack with tensor parallelism on CPU (gloo, 2 ranks): per-rank checkpoint sharding through the builder,
bucketed generation, the speculation verify graph, continuous batching over the page pool and the int8 runner -- each
must reproduce what one process produces."""
from __future__ import annotations

import torch

from conftest import run_distributed


def _reference(sd_seed=0):
    from neuronx_distributed_llama3_2_b200.inference import InferenceConfig, LlamaInferenceModel
    from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM, tiny_llama

    torch.manual_seed(sd_seed)
    sd = {k: v.clone() for k, v in LlamaForCausalLM(tiny_llama(sequence_parallel_enabled=False)).state_dict().items()}
    ref = LlamaInferenceModel(tiny_llama(sequence_parallel_enabled=False),
                              InferenceConfig(batch_size=2, max_context_length=32, seq_len=96))
    ref.load_state_dict(sd)
    return sd, ref


def _tp_worker(rank, world):
    from neuronx_distributed_llama3_2_b200.inference import InferenceConfig, InferenceRunner
    from neuronx_distributed_llama3_2_b200.inference.continuous_batching import ContinuousBatcher
    from neuronx_distributed_llama3_2_b200.models import tiny_llama
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    from neuronx_distributed_llama3_2_b200.inference import _mock_parallel_state

    # the un-sharded oracle: built and run as "tp rank 0 of 1" before the real model-parallel groups exist
    with _mock_parallel_state(1, 0):
        _oracle(rank, world)


def _oracle(rank, world):
    from neuronx_distributed_llama3_2_b200.inference import InferenceConfig, InferenceRunner
    from neuronx_distributed_llama3_2_b200.inference.continuous_batching import ContinuousBatcher
    from neuronx_distributed_llama3_2_b200.models import tiny_llama
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    sd, ref = _reference()
    g = torch.Generator().manual_seed(11)
    prompts = torch.randint(1, 512, (2, 13), generator=g)
    mask = torch.ones_like(prompts)
    mask[1, 8:] = 0
    want = ref.generate(prompts, mask, max_new_tokens=8)
    single = ref.generate(prompts[:1], max_new_tokens=16)
    specs = [(9, 6), (4, 9), (12, 5), (6, 7)]
    cb_prompts = [torch.randint(1, 512, (T,), generator=g) for T, _ in specs]
    cb_want = [ref.generate(p.unsqueeze(0), max_new_tokens=n)[0, p.numel():].tolist() for p, (_, n) in zip(cb_prompts, specs)]
    dsd = {k: v for k, v in sd.items() if "layers.1." not in k}
    ps.set_tensor_model_parallel_size(None)           # end of the mocked single-rank phase
    ps.set_tensor_model_parallel_rank(None)

    ps.initialize_model_parallel(tensor_model_parallel_size=2)
    dev = torch.device("cpu")
    kw = dict(tp_degree=2, max_context_length=32, seq_len=96, buckets=[16, 32], use_cuda_graphs=False)
    runner = InferenceRunner(tiny_llama(sequence_parallel_enabled=False), InferenceConfig(batch_size=2, **kw),
                             checkpoint_loader=lambda: sd, device=dev)
    runner.trace()
    w = runner.model.model.layers[0].mlp.down_proj.weight
    assert w.shape[1] * 2 == sd["model.layers.0.mlp.down_proj.weight"].shape[1]          # really sharded
    torch.testing.assert_close(w, sd["model.layers.0.mlp.down_proj.weight"].chunk(2, 1)[rank])
    got = runner.generate(prompts, mask, max_new_tokens=8)
    assert torch.equal(got, want), (rank, got, want)

    # draft / target speculation, both tensor parallel
    target = InferenceRunner(tiny_llama(sequence_parallel_enabled=False), InferenceConfig(batch_size=1, speculation_length=3, **kw),
                             checkpoint_loader=lambda: sd, device=dev)
    draft = InferenceRunner(tiny_llama(sequence_parallel_enabled=False, num_hidden_layers=1), InferenceConfig(batch_size=1, **kw),
                            checkpoint_loader=lambda: dsd, device=dev)
    target.trace()
    draft.trace()
    out = target.generate_speculative(draft, prompts[:1], max_new_tokens=16)
    assert torch.equal(out, single), (rank, out, single)

    # continuous batching over a small page pool
    ic = InferenceConfig(batch_size=2, tp_degree=2, max_context_length=32, seq_len=96, paged_kv=True, page_size=8,
                         num_kv_pages=5, is_continuous_batching=True)
    cb = InferenceRunner(tiny_llama(sequence_parallel_enabled=False), ic, checkpoint_loader=lambda: sd, device=dev)
    cb.trace()
    engine = ContinuousBatcher(cb.model, max_running=2)
    rids = [engine.submit(p, n) for p, (_, n) in zip(cb_prompts, specs)]
    res = engine.run()
    for rid, wnt in zip(rids, cb_want):
        assert res[rid] == wnt, (rank, rid, res[rid], wnt)

    # int8 weight-only decoder linears, sharded scales: the two ranks agree and stay close to the float logits
    q = InferenceRunner(tiny_llama(sequence_parallel_enabled=False),
                        InferenceConfig(batch_size=2, quantized=True, quantization_type="per_channel_symmetric", **kw),
                        checkpoint_loader=lambda: sd, device=dev)
    q.trace()
    ids = prompts[:, :8]
    pos = torch.arange(8).unsqueeze(0).expand(2, 8)
    _, ql = q.model.forward_infer(ids, pos, return_logits=True)
    _, fl = runner.model.forward_infer(ids, pos, return_logits=True)
    assert ((ql - fl).norm() / fl.norm()).item() < 0.05
    both = [None, None]
    torch.distributed.all_gather_object(both, ql)
    assert torch.equal(both[0], both[1])


def test_tp2_inference_stack_matches_single_process():
    run_distributed(_tp_worker, 2)


def _tp_moe_worker(rank, world):
    from neuronx_distributed_llama3_2_b200.inference import (DbrxInferenceModel, DbrxRunner, InferenceConfig,
                                                             MixtralInferenceModel, MixtralRunner)
    from neuronx_distributed_llama3_2_b200.models import DbrxForCausalLM, MixtralForCausalLM, tiny_dbrx, tiny_mixtral
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps
    from neuronx_distributed_llama3_2_b200.scripts.checkpoint_converter_moe import mock_parallel_state

    families = (("mixtral", MixtralForCausalLM, MixtralInferenceModel, MixtralRunner, tiny_mixtral),
                ("dbrx", DbrxForCausalLM, DbrxInferenceModel, DbrxRunner, tiny_dbrx))
    oracle = {}
    with mock_parallel_state(1, 0):                      # single-rank oracle before the real groups exist
        for name, train_cls, inf_cls, _, cfg_fn in families:
            torch.manual_seed(0)
            mk = lambda: cfg_fn(sequence_parallel_enabled=False, capacity_factor=None)   # noqa: E731
            sd = {k: v.clone() for k, v in train_cls(mk()).state_dict().items()}
            ref = inf_cls(mk(), InferenceConfig(batch_size=2, max_context_length=32, seq_len=64))
            ref.load_state_dict(sd)
            prompts = torch.randint(1, 512, (2, 11), generator=torch.Generator().manual_seed(2))
            oracle[name] = (sd, prompts, ref.generate(prompts, max_new_tokens=6))
    ps.initialize_model_parallel(tensor_model_parallel_size=2)
    for name, _, _, runner_cls, cfg_fn in families:
        sd, prompts, want = oracle[name]
        ic = InferenceConfig(tp_degree=2, batch_size=2, max_context_length=32, seq_len=64, buckets=[16, 32],
                             use_cuda_graphs=False)
        r = runner_cls(cfg_fn(sequence_parallel_enabled=False, capacity_factor=None), ic, checkpoint_loader=lambda: sd,
                       device=torch.device("cpu"))
        r.trace()
        experts = r.model.layers[0].mlp.expert_mlps.mlp_op.down_proj.weight
        assert experts.shape[1] * 2 == sd["layers.0.mlp.expert_mlps.mlp_op.down_proj.weight"].shape[1]     # TP-sharded slabs
        got = r.generate(prompts, max_new_tokens=6)
        assert torch.equal(got, want), (rank, name, got, want)


def test_tp2_moe_inference_matches_single_process():
    run_distributed(_tp_moe_worker, 2)


def _tp4_kv_replication_worker(rank, world):
    """tp = 4 with 2 KV heads: the un-sharded checkpoint goes through ``LlamaAttention.preshard_hook`` (query heads and
    o_proj columns re-ordered next to the replicated KV head of each rank) before it is cut."""
    from neuronx_distributed_llama3_2_b200.inference import InferenceConfig, InferenceRunner, _mock_parallel_state
    from neuronx_distributed_llama3_2_b200.models import tiny_llama
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    with _mock_parallel_state(1, 0):
        sd, ref = _reference()
        prompts = torch.randint(1, 512, (2, 11), generator=torch.Generator().manual_seed(2))
        want = ref.generate(prompts, max_new_tokens=6)
    ps.initialize_model_parallel(tensor_model_parallel_size=4)
    ic = InferenceConfig(tp_degree=4, batch_size=2, max_context_length=32, seq_len=96, buckets=[16, 32], use_cuda_graphs=False)
    r = InferenceRunner(tiny_llama(sequence_parallel_enabled=False), ic, checkpoint_loader=lambda: sd, device=torch.device("cpu"))
    r.trace()
    attn = r.model.model.layers[0].self_attn
    assert attn.kv_size_multiplier == 2 and attn.kv_heads_local == 1 and attn.q_heads_local == 2
    got = r.generate(prompts, max_new_tokens=6)
    assert torch.equal(got, want), (rank, got, want)


def test_tp4_inference_with_replicated_kv_heads():
    run_distributed(_tp4_kv_replication_worker, 4)
