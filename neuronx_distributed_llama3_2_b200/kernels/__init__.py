"""Kernel entry points under the reference's package name (``src/neuronx_distributed/kernels/flash_attn.py``)."""
from ..ops.attention import flash_attn_func, nki_flash_attn_func  # noqa: F401
