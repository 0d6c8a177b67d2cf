"""``neuronx_distributed.utils.medusa_utils`` counterpart (reference ``utils/medusa_utils.py:6-212``): candidate-tree
buffers, candidate generation, greedy posterior evaluation and the input update after an accepted path.  The tree code
lives in ``inference/speculative.py``; this module adds the two list/tensor helpers of the reference module."""
from typing import List, Tuple

import torch

from ..inference.speculative import generate_medusa_buffers  # noqa: F401
from ..inference import speculative as _spec

TOPK = 10


def pad_path(path: List[int], length: int, pad_value: int = -2) -> List[int]:
    """``path`` extended with ``pad_value`` to ``length`` entries (unchanged when already that long)."""
    return list(path) + [pad_value] * max(length - len(path), 0)


# The three functions below keep the reference's calling convention (``utils/medusa_utils.py:120-212``): its graphs sample
# on the device, so what it calls "logits" are already TOKEN IDS -- ``logits [..., k]`` the base model's top-k ids (best
# first), ``medusa_logits [heads, ..., topk]`` every head's top-k ids.  The logit-level variants used by
# ``inference.medusa`` are ``inference.speculative.generate_candidates / evaluate_posterior``.
def generate_candidates(medusa_logits: torch.Tensor, logits: torch.Tensor, tree_indices: torch.Tensor,
                        retrieve_indices: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """``(cartesian candidates [paths, depth + 1], tree candidates [1, nodes])`` from the base model's best token and the
    heads' top-k token ids."""
    root = logits.reshape(-1)[:1]
    flat = torch.cat([root, medusa_logits.reshape(-1)], dim=-1)
    tree = flat[tree_indices]
    ext = torch.cat([tree, torch.zeros(1, dtype=tree.dtype, device=tree.device)], dim=0)      # index -1 -> padding token 0
    return ext[retrieve_indices], tree.unsqueeze(0)


def evaluate_posterior(logits: torch.Tensor, candidates: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Greedy acceptance: ``logits [paths, depth + 1, k]`` holds the verifier's top-k token ids after every candidate
    position; a speculated token is accepted while it equals the verifier's best id one position earlier.  Returns
    ``(best candidate index, accepted length)`` as tensors (index 0 when nothing was accepted)."""
    match = (candidates[:, 1:] == logits[:, :-1, 0].to(candidates.dtype)).int()
    lengths = torch.cumprod(match, dim=1).sum(dim=1)
    accept = lengths.max()
    best = torch.argmax(lengths).to(torch.long) if int(accept) > 0 else torch.zeros((), dtype=torch.long,
                                                                                       device=candidates.device)
    return best, accept


def update_inference_inputs(input_ids: torch.Tensor, candidates: torch.Tensor, best_candidate, accept_length,
                            retrieve_indices: torch.Tensor, outputs, logits: torch.Tensor, medusa_logits: torch.Tensor,
                            new_token: int):
    """Append the accepted prefix of the best path to ``input_ids [1, T]`` and cut ``logits [paths, depth + 1, k]`` /
    ``medusa_logits [heads, paths, depth + 1, topk]`` to the position that predicts the next token.  ``outputs`` (the raw
    model outputs) is accepted for signature parity.  Returns ``(input_ids, logits [1, 1, k], medusa_logits
    [heads, 1, 1, topk], new_token, select_indices)``; ``select_indices`` are the cache positions of the accepted tokens."""
    best, keep = int(best_candidate), int(accept_length) + 1
    prev_len = input_ids.shape[1]
    select_indices = retrieve_indices[best, :keep] + prev_len
    input_ids = torch.cat([input_ids, candidates[best, :keep].reshape(1, keep).to(input_ids.dtype)], dim=-1)
    logits = logits[best, keep - 1:keep].unsqueeze(0)
    medusa_logits = medusa_logits[:, best, keep - 1:keep].unsqueeze(1)
    return input_ids, logits, medusa_logits, new_token + keep, select_indices
