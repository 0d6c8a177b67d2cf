#!/usr/bin/env python
"""Debug: determinism of generate() at batch 12 (batched small-M decode layer), eager and graphs."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from neuronx_distributed_llama3_2_b200.inference import InferenceConfig, InferenceRunner
from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM, tiny_llama

def cfg():
    return tiny_llama(vocab_size=1024, hidden_size=512, intermediate_size=1024, num_hidden_layers=2, num_attention_heads=8,
                      num_key_value_heads=2, head_dim=64, torch_dtype=torch.bfloat16, sequence_parallel_enabled=False)

batch = int(os.environ.get("DBG_BATCH", 12))
torch.manual_seed(1)
dev = torch.device("cuda")
train = LlamaForCausalLM(cfg(), device=dev)
sd = {k: v.detach().cpu().clone() for k, v in train.state_dict().items()}
outs = []
for graphs in (False, True):
    ic = InferenceConfig(batch_size=batch, max_context_length=128, seq_len=256, buckets=[64, 128], max_new_tokens=8,
                         use_cuda_graphs=graphs)
    runner = InferenceRunner(cfg(), ic, checkpoint_loader=lambda: sd, device=dev)
    runner.trace()
    g = torch.Generator(device="cpu").manual_seed(5)
    prompts = torch.randint(1, 1024, (batch, 50), generator=g).to(dev)
    for _ in range(3):
        outs.append(runner.generate(prompts, max_new_tokens=8)[:, -8:].cpu())
names = ["eager1", "eager2", "eager3", "graph1", "graph2", "graph3"]
for i in range(1, 6):
    d = (outs[i] != outs[0])
    print(names[i], "vs eager1: differing tokens", int(d.sum()), "rows", sorted(set(d.nonzero()[:, 0].tolist())), "first step", (d.nonzero()[:, 1].min().item() if d.any() else None))
