from .llama import (LlamaConfig, LlamaForCausalLM, LlamaModel, PRESETS, config_from_hf_json, config_from_json, flops_per_token,  # noqa: F401
                    llama2_7b, llama2_13b, llama2_70b, llama3_1_8b, llama3_1_70b, llama3_2_1b, llama3_2_3b, llama3_8b,
                    llama3_70b, tiny_llama)
from .mixtral import (MixtralConfig, MixtralForCausalLM, convert_mixtral_hf_state_dict,  # noqa: F401
                      convert_mixtral_state_dict_to_hf, mixtral_8x7b, tiny_mixtral)
from .gpt_neox import GPTNeoXConfig, GPTNeoXForCausalLM, codegen25_7b, gpt_neox_20b, tiny_gpt_neox  # noqa: F401
from .bert import BertConfig, BertForPreTraining, bert_large, convert_bert_hf_state_dict, tiny_bert  # noqa: F401
from .dbrx import DbrxConfig, DbrxForCausalLM, convert_dbrx_hf_state_dict, dbrx_instruct, tiny_dbrx  # noqa: F401
