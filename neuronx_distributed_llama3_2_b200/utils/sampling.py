"""``neuronx_distributed.utils.sampling`` counterpart (reference ``utils/sampling.py:6-77``): on-device greedy /
top-k / top-p sampling; the kernels are in ``csrc/sampling.cu``."""
from ..inference.sampling import Sampler  # noqa: F401
