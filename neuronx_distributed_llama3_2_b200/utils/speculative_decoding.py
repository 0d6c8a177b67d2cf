"""``neuronx_distributed.utils.speculative_decoding`` counterpart (reference ``utils/speculative_decoding.py:15-304``)."""
from ..inference.speculative import NeuronSpeculation  # noqa: F401
