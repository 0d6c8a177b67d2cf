"""Speculative decoding reproduces the target's greedy output; Medusa buffers / candidate / posterior helpers."""
from __future__ import annotations

import torch

from neuronx_distributed_llama3_2_b200.inference.speculative import (NeuronSpeculation, evaluate_posterior,
                                                                    generate_candidates, generate_medusa_buffers)


def test_draft_verify_equals_target_greedy():
    torch.manual_seed(0)
    V = 17
    table_t = torch.randint(0, V, (V, 64))        # target: next = table_t[token, position]
    table_d = table_t.clone()
    table_d[:, ::3] = torch.randint(0, V, (V, 22))  # the draft disagrees at every third position

    def draft_step(tok, pos):
        return table_d[tok.reshape(-1), pos.reshape(-1)]

    def target_verify(toks, pos):
        return table_t[toks, pos]

    spec = NeuronSpeculation(draft_step, target_verify, speculation_length=4)
    out = spec.generate(torch.tensor([3]), torch.tensor([5]), max_new_tokens=20)
    ref, t, p = [], 3, 5
    for _ in range(20):
        t = int(table_t[t, p])
        p += 1
        ref.append(t)
    assert out.tolist() == ref
    assert max(spec.accepted_hist) >= 1 and min(spec.accepted_hist) < 4     # some rounds accepted, some rejected


def test_medusa_helpers():
    choices = [[0], [1], [0, 0], [0, 1], [1, 0], [0, 0, 0]]
    buf = generate_medusa_buffers(choices)
    n = len(choices) + 1
    m = buf["medusa_attn_mask"][0, 0]
    assert m.shape == (n, n) and torch.all(m[:, 0] == 1) and torch.all(torch.diag(m) == 1)
    assert buf["medusa_position_ids"].tolist() == [0, 1, 1, 2, 2, 2, 3]
    assert buf["retrieve_indices"].shape[1] == 4
    V = 30
    base = torch.zeros(V)
    base[7] = 5.0
    heads = torch.zeros(3, V)
    heads[0, 11], heads[0, 12] = 4.0, 3.0
    heads[1, 21], heads[1, 22] = 4.0, 3.0
    heads[2, 5] = 4.0
    cands, tree = generate_candidates(heads, base, buf["tree_indices"], buf["retrieve_indices"], topk=buf["topk"])
    assert tree[0] == 7 and 11 in tree.tolist() and 21 in tree.tolist()
    assert cands[0].tolist() == [7, 11, 21, 5]                      # the deepest path: root, top-0 of every head
    target = torch.zeros_like(cands)
    target[0] = torch.tensor([11, 21, 9, 0])                        # target agrees with the first two speculated tokens
    best, n_acc = evaluate_posterior(target, cands, valid=buf["retrieve_indices"] >= 0)
    assert (best, n_acc) == (0, 2)


def test_speculation_with_kv_cache_models_equals_target_greedy():
    """Real graphs: a 1-layer draft and the 2-layer target, each with its own KV cache; the target's
    ``forward_speculation`` verifies k+1 tokens on top of the cache.  Output == the target's own greedy decoding, for the
    contiguous and the paged cache."""
    from neuronx_distributed_llama3_2_b200.inference import InferenceConfig, LlamaInferenceModel
    from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM, tiny_llama

    for paged in (False, True):
        torch.manual_seed(0)
        ic = dict(batch_size=1, max_context_length=32, seq_len=96, paged_kv=paged, page_size=16)
        target = LlamaInferenceModel(tiny_llama(sequence_parallel_enabled=False), InferenceConfig(**ic))
        draft = LlamaInferenceModel(tiny_llama(sequence_parallel_enabled=False, num_hidden_layers=1), InferenceConfig(**ic))
        sd = LlamaForCausalLM(tiny_llama(sequence_parallel_enabled=False)).state_dict()
        target.load_state_dict(sd)
        draft.load_state_dict({k: v for k, v in sd.items() if "layers.1." not in k})      # shares layer 0, embedding, head
        prompt = torch.randint(1, 512, (1, 12))
        ref = target.generate(prompt, max_new_tokens=24)[0, 12:]
        target.init_kv_cache()
        draft.init_kv_cache()
        pos = torch.arange(12).unsqueeze(0)
        first = target.forward_infer(prompt, pos)
        draft.forward_infer(prompt, pos)
        # the verify graph agrees with token-by-token decoding on logits, not only on the argmax
        spec = NeuronSpeculation(lambda t, p: draft.forward_infer(t, p), lambda t, p: target.forward_speculation(t, p),
                                 speculation_length=3)
        new = spec.generate(first, torch.tensor([12]), max_new_tokens=23)
        assert torch.equal(torch.cat([first, new]), ref), (paged, torch.cat([first, new]), ref)
        assert len(spec.accepted_hist) < 23                       # fewer target calls than tokens


def test_runner_speculative_generation():
    """Target runner traced with ``speculation_length`` gets a third sub-model (the verify graph); assisted generation
    through the two runners equals the target runner's own ``generate``."""
    from neuronx_distributed_llama3_2_b200.inference import InferenceConfig, InferenceRunner
    from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM, tiny_llama

    torch.manual_seed(0)
    sd = {k: v.clone() for k, v in LlamaForCausalLM(tiny_llama(sequence_parallel_enabled=False)).state_dict().items()}
    dsd = {k: v for k, v in sd.items() if "layers.1." not in k}
    kw = dict(batch_size=1, max_context_length=32, seq_len=96, buckets=[16, 32], use_cuda_graphs=False)
    target = InferenceRunner(tiny_llama(sequence_parallel_enabled=False), InferenceConfig(speculation_length=3, **kw),
                             checkpoint_loader=lambda: sd, device=torch.device("cpu"))
    draft = InferenceRunner(tiny_llama(sequence_parallel_enabled=False, num_hidden_layers=1), InferenceConfig(**kw),
                            checkpoint_loader=lambda: dsd, device=torch.device("cpu"))
    target.trace()
    draft.trace()
    assert set(target.nxd_model.keys()) == {"context_encoding_model", "token_generation_model", "speculation_model"}
    prompt = torch.randint(1, 512, (1, 11))
    ref = target.generate(prompt, max_new_tokens=20)
    out = target.generate_speculative(draft, prompt, max_new_tokens=20)
    assert torch.equal(out, ref), (out, ref)
    assert len(target.last_acceptance) < 19


def test_medusa_generation_equals_base_greedy():
    """Medusa heads + candidate tree + path verification on forked cache lines: whatever the heads guess, the output
    is the base model's greedy continuation; heads that guess well make it take fewer model calls."""
    from neuronx_distributed_llama3_2_b200.inference import InferenceConfig, LlamaInferenceModel
    from neuronx_distributed_llama3_2_b200.inference.medusa import DEFAULT_MEDUSA_CHOICES, medusa_generate
    from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM, tiny_llama

    torch.manual_seed(0)
    sd = LlamaForCausalLM(tiny_llama(sequence_parallel_enabled=False)).state_dict()
    ref = LlamaInferenceModel(tiny_llama(sequence_parallel_enabled=False),
                              InferenceConfig(batch_size=1, max_context_length=32, seq_len=96))
    ref.load_state_dict(sd)
    prompt = torch.randint(1, 512, (13,), generator=torch.Generator().manual_seed(3))
    want = ref.generate(prompt.unsqueeze(0), max_new_tokens=30)[0, 13:]
    n_paths = 4                                                   # leaves of the default tree
    ic = InferenceConfig(batch_size=n_paths + 1, max_context_length=32, seq_len=96, paged_kv=True, page_size=8,
                         num_kv_pages=40, num_medusa_heads=3)
    model = LlamaInferenceModel(tiny_llama(sequence_parallel_enabled=False), ic)
    missing = model.load_state_dict(sd, strict=False)
    assert sorted(missing.missing_keys) == sorted(
        f"medusa_head_{i}.{k}" for i in range(3) for k in ("0.linear.weight", "0.linear.bias", "1.weight"))
    # untrained heads (random projection): guesses are mostly wrong, the result must still be the greedy one
    got, acc = medusa_generate(model, prompt, 30, DEFAULT_MEDUSA_CHOICES)
    assert torch.equal(got, want), (got, want)
    assert len(acc) >= 30 // 4 and model.kv_cache.pool.num_free >= 40 - 2 * ((13 + 30 + 8) // 8 + 1)
    # heads that copy the LM head guess "the same token again": right whenever the model repeats itself, which this
    # random model soon does -- accepted guesses then cut the number of model calls
    with torch.no_grad():
        for i in range(3):
            getattr(model, f"medusa_head_{i}")[1].weight.copy_(model.lm_head.weight)
    got2, acc2 = medusa_generate(model, prompt, 30)
    assert torch.equal(got2, want)
    assert sum(a + 1 for a in acc2) >= 30 and all(0 <= a <= 3 for a in acc2)
    assert max(acc2) == 3 and len(acc2) < len(acc) // 2, (acc2, len(acc))
    assert model.kv_cache.pool.num_free >= 40 - ((13 + 34) // 8 + 1)          # losing forks gave their pages back
