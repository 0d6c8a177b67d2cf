"""Medusa decoding on the KV-cache inference models (reference ``examples/inference/llama3/
neuron_modeling_llama.py:416-430`` -- ``medusa_head_i = Sequential(ResBlock, lm-head-shaped linear)`` -- and the Medusa
branches of ``modules/model_base.py`` driven by ``utils/medusa_utils.py``).

The extra heads read the final-norm hidden state of the last accepted token; head ``i`` guesses the token ``i + 2``
positions ahead.  Their top-k guesses are arranged into a candidate tree (``generate_medusa_buffers`` /
``generate_candidates``), the base model scores the candidates, and the longest prefix that matches the base model's own
greedy choices is accepted (``evaluate_posterior``) -- so the output is exactly the base model's greedy continuation,
produced in fewer model calls.

Verification: the reference evaluates all tree nodes in one pass under a tree attention mask.  Here every root-to-leaf
PATH is verified as its own short sequence by the ``speculation_model`` graph (``forward_speculation``): the accepted
prefix lives on one cache line, each path gets a copy-on-write FORK of that line from the page pool (full pages shared,
only the partial page copied), the winning fork becomes the new main line.  Same tokens as tree attention, more rows
(paths x depth instead of tree nodes) -- the price of reusing the decode-attention kernel unchanged.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch
import torch.nn as nn

from ..parallel import mappings
from ..parallel.layers import ColumnParallelLinear
from .speculative import evaluate_posterior, generate_candidates, generate_medusa_buffers

# a small default tree: 2 guesses of head 0, the better of them continued by head 1 and head 2
DEFAULT_MEDUSA_CHOICES = [[0], [1], [0, 0], [0, 1], [1, 0], [0, 0, 0]]


class ResBlock(nn.Module):
    """``x + SiLU(Linear(x))`` with the linear zero-initialised, so a fresh head starts as a copy of its input
    (reference ``neuron_modeling_llama.py`` ``ResBlock``)."""

    def __init__(self, hidden_size: int, dtype=None, device=None):
        super().__init__()
        self.linear = nn.Linear(hidden_size, hidden_size, dtype=dtype, device=device)
        if self.linear.weight.device.type != "meta":
            nn.init.zeros_(self.linear.weight)
            nn.init.zeros_(self.linear.bias)
        self.act = nn.SiLU()

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return x + self.act(self.linear(x))


def add_medusa_heads(model, num_heads: int, device=None) -> None:
    """Attach ``medusa_head_0 .. medusa_head_{n-1}`` (parameter names as the reference: ``medusa_head_i.0.linear.*``,
    ``medusa_head_i.1.weight``) to a KV-cache inference model.  The vocabulary projection is column parallel like the LM
    head."""
    cfg = model.config
    for i in range(num_heads):
        head = nn.Sequential(
            ResBlock(cfg.hidden_size, dtype=cfg.torch_dtype, device=device),
            ColumnParallelLinear(cfg.hidden_size, cfg.vocab_size, bias=False, gather_output=False, dtype=cfg.torch_dtype,
                                 device=device))
        setattr(model, f"medusa_head_{i}", head)
    model.num_medusa_heads = num_heads


def medusa_logits(model, hidden: torch.Tensor) -> torch.Tensor:
    """``hidden [..., H]`` (final-norm output) -> ``[heads, ..., V]`` full-vocabulary logits of every Medusa head."""
    outs = []
    for i in range(model.num_medusa_heads):
        lg = getattr(model, f"medusa_head_{i}")(hidden)
        lg = mappings.gather_from_tensor_model_parallel_region(lg)[..., : model.config.vocab_size]
        outs.append(lg)
    return torch.stack(outs, dim=0)


@torch.no_grad()
def medusa_generate(model, prompt: torch.Tensor, max_new_tokens: int, medusa_choices: Optional[List[List[int]]] = None,
                    eos_token_id: Optional[int] = None) -> Tuple[torch.Tensor, List[int]]:
    """Greedy Medusa decoding of ONE prompt ``[T]``.  ``model`` needs Medusa heads (``InferenceConfig.num_medusa_heads``),
    the dynamic page pool (``paged_kv=True, num_kv_pages=N``) and at least ``paths + 1`` cache lines.  Returns
    ``(generated tokens [<= max_new_tokens], accepted-speculated-token count of every round)``."""
    cache = model.kv_cache if model.kv_cache is not None else model.init_kv_cache()
    assert cache.pool is not None, "medusa_generate forks cache lines: InferenceConfig(paged_kv=True, num_kv_pages=N)"
    dev = cache.block_table.device
    buf = generate_medusa_buffers(medusa_choices or DEFAULT_MEDUSA_CHOICES, device=dev)
    retrieve, tree_indices, topk = buf["retrieve_indices"], buf["tree_indices"], buf["topk"]
    n_paths, depth1 = retrieve.shape
    assert int(retrieve.max()) < tree_indices.numel()
    assert model.num_medusa_heads >= depth1 - 1, f"the tree is {depth1 - 1} deep but there are {model.num_medusa_heads} heads"
    assert cache.lines >= n_paths + 1, f"{n_paths} candidate paths need {n_paths + 1} cache lines, the cache has {cache.lines}"
    valid = retrieve >= 0
    main = 0
    scratch = list(range(1, n_paths + 1))
    for line in range(cache.lines):
        cache.free(line)
    prompt = prompt.reshape(1, -1).to(dev)
    T = prompt.shape[1]
    cache.reserve(main, T)
    pos = torch.arange(T, device=dev).unsqueeze(0)
    _, base, hidden = model.forward_infer(prompt, pos, seq_ids=torch.tensor([main], dtype=torch.int32, device=dev),
                                          last_token_index=torch.tensor([T - 1], device=dev), return_hidden=True)
    base, heads = base[0], medusa_logits(model, hidden)[:, 0]             # [V], [heads, V]
    out: List[int] = []
    accepted: List[int] = []
    cur = T                                                              # tokens of the accepted prefix in the cache
    while len(out) < max_new_tokens:
        cands, _ = generate_candidates(heads[: depth1 - 1], base, tree_indices, retrieve, topk=topk)   # [paths, depth+1]
        cache.truncate(main, cur)                                        # the allocator's length = the accepted prefix
        for line in scratch:                                             # every path continues from the accepted prefix
            cache.fork(main, line)
            cache.reserve(line, cur + depth1)
        lines = torch.tensor(scratch, dtype=torch.int32, device=dev)
        ppos = (cur + torch.arange(depth1, device=dev)).unsqueeze(0).expand(n_paths, depth1).contiguous()
        target, logits, hidden = model.forward_speculation(cands.contiguous(), ppos, seq_ids=lines, return_logits=True,
                                                           return_hidden=True)
        best, n_acc = evaluate_posterior(target, cands, valid=valid)
        new = cands[best, : n_acc + 1].tolist()                          # the root (always right) + the accepted guesses
        accepted.append(n_acc)
        # the winning fork holds the K/V of the accepted tokens: it simply becomes the main line (no copy)
        losers = [line for line in scratch if line != scratch[best]] + [main]
        main = scratch[best]
        for line in losers:
            cache.free(line)
        scratch = losers
        cur += len(new)
        base, heads = logits[best, n_acc], medusa_logits(model, hidden[best, n_acc])
        for tok in new:
            out.append(tok)
            if eos_token_id is not None and tok == eos_token_id:
                return torch.tensor(out[:max_new_tokens], device=dev), accepted
    return torch.tensor(out[:max_new_tokens], device=dev), accepted


__all__ = ["DEFAULT_MEDUSA_CHOICES", "ResBlock", "add_medusa_heads", "medusa_generate", "medusa_logits"]
