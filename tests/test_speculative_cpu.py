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
