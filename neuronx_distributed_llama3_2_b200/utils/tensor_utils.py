"""``neuronx_distributed.utils.tensor_utils`` counterpart (reference ``utils/tensor_utils.py:4-62``)."""
import torch


def cumsum(tensor: torch.Tensor, dim: int = 0, tril_size: int = 2048) -> torch.Tensor:
    """Cumulative sum along ``dim``.  The reference multiplies by lower-triangular blocks of ``tril_size`` because the
    XLA cumsum lowering is slow on its hardware; on CUDA ``torch.cumsum`` is a single scan kernel, so ``tril_size`` is
    accepted for signature compatibility only.  Integer inputs are accumulated exactly (the reference goes through
    floating-point matmuls)."""
    del tril_size
    return torch.cumsum(tensor, dim=dim)
