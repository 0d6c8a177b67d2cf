"""Speculative decoding: draft-k-then-verify (greedy) and Medusa tree decoding helpers (reference
``src/neuronx_distributed/utils/speculative_decoding.py:40-304`` and ``utils/medusa_utils.py:32-212``)."""
from __future__ import annotations

from typing import Callable, List, Optional, Tuple

import torch


class NeuronSpeculation:
    """Assisted (draft/target) greedy decoding.

    ``draft_step(tokens [B,1], pos [B,1]) -> next token [B]`` and ``target_verify(tokens [B,k+1], pos [B,k+1]) ->
    greedy target tokens [B,k+1]`` are the two compiled graphs (the reference's ``speculation_model`` is the target
    run on ``speculation_length`` tokens at once, ``model_base.py:853-860``).  Each round proposes ``k`` draft tokens,
    verifies them in ONE target call and keeps the longest matching prefix plus the target's correction token, so the
    output is exactly what the target model alone would produce.
    """

    def __init__(self, draft_step: Callable, target_verify: Callable, speculation_length: int = 4):
        assert speculation_length >= 1
        self.draft_step, self.target_verify, self.k = draft_step, target_verify, speculation_length
        self.accepted_hist: List[int] = []

    @torch.no_grad()
    def generate(self, last_token: torch.Tensor, position: torch.Tensor, max_new_tokens: int,
                 eos_token_id: Optional[int] = None) -> torch.Tensor:
        """``last_token [1]`` = last accepted token (already in both KV caches up to ``position - 1``),
        ``position [1]`` = its position.  Batch size 1 (as the reference).  Returns the new tokens ``[n]``."""
        out: List[int] = []
        tok, pos = last_token.reshape(1, 1), position.reshape(1, 1)
        while len(out) < max_new_tokens:
            drafts = []
            t, p = tok, pos
            for _ in range(self.k):
                t = self.draft_step(t, p).reshape(1, 1)
                p = p + 1
                drafts.append(t)
            cand = torch.cat([tok] + drafts, dim=1)                                  # [1, k+1]
            cand_pos = pos + torch.arange(self.k + 1, device=pos.device).unsqueeze(0)
            target = self.target_verify(cand, cand_pos).reshape(1, self.k + 1)       # target's next token at each slot
            draft_tokens = cand[:, 1:]
            match = (draft_tokens == target[:, : self.k]).long().cumprod(dim=1)
            n_ok = int(match.sum())
            self.accepted_hist.append(n_ok)
            new = draft_tokens[0, :n_ok].tolist() + [int(target[0, n_ok])]
            for x in new:
                out.append(x)
                if eos_token_id is not None and x == eos_token_id:
                    return torch.tensor(out[:max_new_tokens], device=last_token.device)
            if n_ok == self.k:
                # every draft accepted: the draft model never consumed its own last proposal, so its KV cache has a hole
                # at that position -- fill it (the prediction is discarded) or the next round drafts from a stale context
                self.draft_step(drafts[-1], pos + self.k)
            tok = torch.tensor([[new[-1]]], device=last_token.device)
            pos = pos + len(new)
        return torch.tensor(out[:max_new_tokens], device=last_token.device)


# ---------------------------------------------------------------------------------------------------------------------
# Medusa
# ---------------------------------------------------------------------------------------------------------------------

def generate_medusa_buffers(medusa_choices: List[List[int]], device="cpu") -> dict:
    """Static buffers describing the candidate tree (reference ``medusa_utils.py:32-118``).

    ``medusa_choices`` lists tree paths; path ``[a, b]`` means "the a-th top prediction of head 0 followed by the b-th
    top prediction of head 1".  Returns the tree attention mask (ancestor visibility incl. the root), the index of
    every tree node in the flattened candidate list, per-node depth (position offsets) and, for every leaf-to-root
    path, the node indices used to read back complete candidate sequences."""
    paths = sorted(medusa_choices, key=lambda x: (len(x), x))
    n = len(paths) + 1
    depth_counts: List[int] = []
    for p in paths:
        d = len(p)
        while len(depth_counts) < d:
            depth_counts.append(0)
        depth_counts[d - 1] += 1
    mask = torch.eye(n)
    mask[:, 0] = 1.0
    index_of = {tuple(p): i + 1 for i, p in enumerate(paths)}
    for p in paths:
        i = index_of[tuple(p)]
        for c in range(1, len(p)):
            mask[i, index_of[tuple(p[:c])]] = 1.0
    topk = max(max(p) for p in paths) + 1
    tree_indices = torch.zeros(n, dtype=torch.long)
    for p in paths:
        tree_indices[index_of[tuple(p)]] = p[-1] + topk * (len(p) - 1) + 1
    position_ids = torch.zeros(n, dtype=torch.long)
    for p in paths:
        position_ids[index_of[tuple(p)]] = len(p)
    leaves = [p for p in paths if not any(len(q) > len(p) and q[: len(p)] == p for q in paths)]
    max_len = max(len(p) for p in leaves) + 1
    retrieve = torch.full((len(leaves), max_len), -1, dtype=torch.long)
    for r, p in enumerate(sorted(leaves, key=lambda x: (-len(x), x))):
        retrieve[r, 0] = 0
        for c in range(len(p)):
            retrieve[r, c + 1] = index_of[tuple(p[: c + 1])]
    return {"medusa_attn_mask": mask.unsqueeze(0).unsqueeze(0).to(device), "tree_indices": tree_indices.to(device),
            "medusa_position_ids": position_ids.to(device), "retrieve_indices": retrieve.to(device), "topk": topk}


def generate_candidates(medusa_logits: torch.Tensor, base_logits: torch.Tensor, tree_indices: torch.Tensor,
                        retrieve_indices: torch.Tensor, topk: int = 10) -> Tuple[torch.Tensor, torch.Tensor]:
    """``base_logits [V]`` (target head, last position), ``medusa_logits [heads, V]`` -> ``(candidates [paths, depth+1],
    tree_candidates [nodes])`` (reference ``medusa_utils.py:121-166``)."""
    root = torch.argmax(base_logits, dim=-1).reshape(1)
    heads_topk = torch.topk(medusa_logits, topk, dim=-1).indices.reshape(-1)
    flat = torch.cat([root, heads_topk])
    tree_candidates = flat[tree_indices]
    ext = torch.cat([tree_candidates, torch.zeros(1, dtype=torch.long, device=flat.device)])
    cands = ext[retrieve_indices]           # -1 picks the appended zero
    return cands, tree_candidates


def evaluate_posterior(target_tokens: torch.Tensor, candidates: torch.Tensor, valid: Optional[torch.Tensor] = None
                       ) -> Tuple[int, int]:
    """Greedy acceptance (reference ``medusa_utils.py:169-212``): ``target_tokens [paths, depth+1]`` = the target's
    argmax after each candidate position.  Candidate token c+1 is accepted iff it equals the target's prediction
    after token c.  Returns ``(best path index, number of accepted speculated tokens)``."""
    ok = (candidates[:, 1:] == target_tokens[:, :-1])
    if valid is not None:
        ok = ok & valid[:, 1:]
    acc = ok.long().cumprod(dim=1).sum(dim=1)
    best = int(torch.argmax(acc))
    return best, int(acc[best])
