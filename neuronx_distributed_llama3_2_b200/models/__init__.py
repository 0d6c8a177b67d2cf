from .llama import (LlamaConfig, LlamaForCausalLM, LlamaModel, PRESETS, flops_per_token, llama3_1_8b,  # noqa: F401
                    llama3_2_1b, llama3_2_3b, tiny_llama)
from .mixtral import MixtralConfig, MixtralForCausalLM, mixtral_8x7b, tiny_mixtral  # noqa: F401
