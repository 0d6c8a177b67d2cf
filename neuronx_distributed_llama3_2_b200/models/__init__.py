from .llama import (LlamaConfig, LlamaForCausalLM, LlamaModel, PRESETS, flops_per_token, llama3_1_8b,  # noqa: F401
                    llama3_2_1b, llama3_2_3b, tiny_llama)
from .mixtral import MixtralConfig, MixtralForCausalLM, mixtral_8x7b, tiny_mixtral  # noqa: F401
from .gpt_neox import GPTNeoXConfig, GPTNeoXForCausalLM, codegen25_7b, gpt_neox_20b, tiny_gpt_neox  # noqa: F401
from .bert import BertConfig, BertForPreTraining, bert_large, tiny_bert  # noqa: F401
from .dbrx import DbrxConfig, DbrxForCausalLM, convert_dbrx_hf_state_dict, dbrx_instruct, tiny_dbrx  # noqa: F401
