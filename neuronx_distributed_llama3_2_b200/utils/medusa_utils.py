"""``neuronx_distributed.utils.medusa_utils`` counterpart (reference ``utils/medusa_utils.py:6-212``): candidate-tree
buffers, candidate generation, greedy posterior evaluation and the input update after an accepted path.  The tree code
lives in ``inference/speculative.py``; this module adds the two list/tensor helpers of the reference module."""
from typing import List, Tuple

import torch

from ..inference.speculative import evaluate_posterior, generate_candidates, generate_medusa_buffers  # noqa: F401

TOPK = 10


def pad_path(path: List[int], length: int, pad_value: int = -2) -> List[int]:
    """``path`` extended with ``pad_value`` to ``length`` entries (unchanged when already that long)."""
    return list(path) + [pad_value] * max(length - len(path), 0)


def update_inference_inputs(input_ids: torch.Tensor, candidates: torch.Tensor, best_candidate: int, accept_length: int,
                            retrieve_indices: torch.Tensor, logits: torch.Tensor, medusa_logits: torch.Tensor,
                            new_token: int) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor, int, torch.Tensor]:
    """Append the accepted prefix of the best candidate path to ``input_ids [1, T]`` and cut the logits to the position
    that predicts the next token.  ``logits [paths, depth+1, V]``, ``medusa_logits [heads, paths, depth+1, V]``.
    Returns ``(input_ids, logits [1, 1, V], medusa_logits [heads, 1, 1, V], new_token, select_indices)`` where
    ``select_indices`` are the KV-cache rows (tree-node positions shifted by the old length) of the accepted tokens."""
    keep = int(accept_length) + 1
    prev_len = input_ids.shape[1]
    select_indices = retrieve_indices[best_candidate, :keep] + prev_len
    input_ids = torch.cat([input_ids, candidates[best_candidate, :keep].reshape(1, keep).to(input_ids.dtype)], dim=-1)
    logits = logits[best_candidate, keep - 1:keep].unsqueeze(0)
    medusa_logits = medusa_logits[:, best_candidate, keep - 1:keep].unsqueeze(1)
    return input_ids, logits, medusa_logits, new_token + keep, select_indices
