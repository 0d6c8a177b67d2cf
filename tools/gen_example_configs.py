#!/usr/bin/env python
"""Write ``examples/training/configs/<name>/config.json`` (HF ``config.json`` layout) from the presets in ``models/`` -- the
counterpart of the per-size config directories the reference ships next to its launch scripts
(``examples/training/llama/tp_zero1_llama_hf_pretrain/*_config_*``, ``tp_pp_llama_hf_pretrain/*_config_*``,
``mixtral_pretrain/configs``, ``codegen25/config.json``).  Re-run after changing a preset."""
from __future__ import annotations

import json
import os
import sys
from dataclasses import asdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from neuronx_distributed_llama3_2_b200 import models as M  # noqa: E402

RUNTIME_KEYS = ("sequence_parallel_enabled", "kv_shared_group_size", "fuse_qkv", "use_flash_attention",
                "selective_checkpoint_enabled", "activation_checkpointing", "extra")


def dump(name: str, cfg, architectures):
    d = asdict(cfg)
    for k in RUNTIME_KEYS:
        d.pop(k, None)
    d["torch_dtype"] = str(cfg.torch_dtype).replace("torch.", "")
    d["architectures"] = architectures
    out = os.path.join(ROOT, "examples", "training", "configs", name)
    os.makedirs(out, exist_ok=True)
    with open(os.path.join(out, "config.json"), "w") as f:
        json.dump(d, f, indent=2, sort_keys=True)
        f.write("\n")


if __name__ == "__main__":
    for key, fn in M.PRESETS.items():
        if key != "tiny":
            dump(key.replace("-", "_"), fn(), ["LlamaForCausalLM"])
    dump("mixtral_8x7b", M.mixtral_8x7b(), ["MixtralForCausalLM"])
    dump("dbrx_instruct", M.dbrx_instruct(), ["DbrxForCausalLM"])
    dump("gpt_neox_20b", M.gpt_neox_20b(), ["GPTNeoXForCausalLM"])
    dump("gpt_neox_6.9b", M.gpt_neox_20b(hidden_size=4096, num_hidden_layers=32, num_attention_heads=32,
                                         intermediate_size=16384), ["GPTNeoXForCausalLM"])
    dump("codegen25_7b", M.codegen25_7b(), ["CodeGenForCausalLM"])
    dump("bert_large", M.bert_large(), ["BertForPreTraining"])
    print(sorted(os.listdir(os.path.join(ROOT, "examples", "training", "configs"))))
