"""GPT-NeoX / CodeGen-style and BERT models: TP = 2 (+ sequence parallelism) against the single-process model with
the same weights -- loss and gradients (reference: the example training scripts of these families only run end to end;
the parity harness is the one of ``test/integration/parallel_layers/test_layers.py``)."""
from __future__ import annotations

import contextlib

import pytest
import torch

from conftest import run_distributed


@contextlib.contextmanager
def _solo():
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    ps.set_tensor_model_parallel_size(1)
    ps.set_tensor_model_parallel_rank(0)
    try:
        yield
    finally:
        ps.set_tensor_model_parallel_size(None)
        ps.set_tensor_model_parallel_rank(None)


def _local(w, p, tp, rank):
    from neuronx_distributed_llama3_2_b200.parallel.layers import create_local_weight

    if not getattr(p, "tensor_model_parallel", False):
        return w
    return create_local_weight(w, p.partition_dim, w.shape[p.partition_dim] // tp, p.partition_stride, rank=rank,
                               world_size=tp)


def _build(family, sp):
    if family == "bert":
        from neuronx_distributed_llama3_2_b200.models.bert import BertForPreTraining, tiny_bert

        return BertForPreTraining(tiny_bert(sequence_parallel_enabled=sp))
    from neuronx_distributed_llama3_2_b200.models.gpt_neox import GPTNeoXForCausalLM, tiny_gpt_neox

    kw = {} if family == "neox" else dict(rotary_interleaved=True, shared_layernorm=True, qkv_bias=False, rotary_pct=0.5,
                                          hidden_act="gelu_new")
    return GPTNeoXForCausalLM(tiny_gpt_neox(sequence_parallel_enabled=sp, **kw))


def _batch(family):
    g = torch.Generator().manual_seed(3)
    ids = torch.randint(0, 512, (2, 32), generator=g)
    if family != "bert":
        return dict(input_ids=ids, labels=ids)
    labels = torch.full_like(ids, -100)
    pick = torch.rand(ids.shape, generator=g) < 0.3
    labels[pick] = ids[pick]
    types = (torch.arange(32) >= 16).long().unsqueeze(0).expand(2, 32).contiguous()
    return dict(input_ids=ids, token_type_ids=types, labels=labels, next_sentence_label=torch.tensor([0, 1]))


def _worker(rank, world, family, sp):
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps
    from neuronx_distributed_llama3_2_b200.parallel.grads import allreduce_sequence_parallel_gradients

    torch.manual_seed(0)
    with _solo():
        full = _build(family, False)
    batch = _batch(family)
    with _solo():
        ref = full(**batch)
        ref.backward()
    ps.initialize_model_parallel(tensor_model_parallel_size=world)
    model = _build(family, sp)
    full_p = dict(full.named_parameters())
    with torch.no_grad():
        for n, p in model.named_parameters():
            p.copy_(_local(full_p[n], p, world, rank))
    loss = model(**batch)
    loss.backward()
    assert abs(loss.item() - ref.item()) < 2e-5, (family, sp, loss.item(), ref.item())
    if sp:
        allreduce_sequence_parallel_gradients(list(model.parameters()))
    checked = 0
    for n, p in model.named_parameters():
        want = _local(full_p[n].grad, p, world, rank)
        torch.testing.assert_close(p.grad, want, atol=3e-5, rtol=2e-3, msg=lambda m: f"{family} sp={sp} {n}: {m}")
        checked += 1
    assert checked >= 15
    ps.destroy_model_parallel()


@pytest.mark.parametrize("family,sp", [("neox", True), ("neox", False), ("codegen", True), ("bert", True), ("bert", False)])
def test_family_tp2_matches_single_process(family, sp):
    run_distributed(_worker, 2, family, sp)


def test_bert_padding_mask_and_logits_single_process():
    from neuronx_distributed_llama3_2_b200.models.bert import BertForPreTraining, tiny_bert

    torch.manual_seed(1)
    m = BertForPreTraining(tiny_bert()).eval()
    ids = torch.randint(0, 512, (2, 24))
    mask = torch.ones(2, 24, dtype=torch.long)
    mask[1, 16:] = 0
    with torch.no_grad():
        logits, nsp = m(ids, attention_mask=mask)
        # tokens behind the padding must not influence the real positions of that row
        ids2 = ids.clone()
        ids2[1, 16:] = 7
        logits2, nsp2 = m(ids2, attention_mask=mask)
    assert logits.shape == (24, 2, 512) and nsp.shape == (2, 2)
    torch.testing.assert_close(logits[:16, 1], logits2[:16, 1], atol=1e-5, rtol=1e-5)
    torch.testing.assert_close(nsp, nsp2, atol=1e-5, rtol=1e-5)
