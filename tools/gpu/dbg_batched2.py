#!/usr/bin/env python
"""Debug: KV cache contents after prefill + N decode steps, eager vs graphs, batch 12."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from neuronx_distributed_llama3_2_b200.inference import InferenceConfig, InferenceRunner
from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM, tiny_llama

def cfg():
    return tiny_llama(vocab_size=1024, hidden_size=512, intermediate_size=1024, num_hidden_layers=2, num_attention_heads=8,
                      num_key_value_heads=2, head_dim=64, torch_dtype=torch.bfloat16, sequence_parallel_enabled=False)

batch = int(os.environ.get("DBG_BATCH", 12))
new = int(os.environ.get("DBG_NEW", 3))
torch.manual_seed(1)
dev = torch.device("cuda")
train = LlamaForCausalLM(cfg(), device=dev)
sd = {k: v.detach().cpu().clone() for k, v in train.state_dict().items()}
caches, toks = [], []
for graphs in (False, True):
    ic = InferenceConfig(batch_size=batch, max_context_length=128, seq_len=256, buckets=[64, 128], max_new_tokens=8,
                         use_cuda_graphs=graphs)
    runner = InferenceRunner(cfg(), ic, checkpoint_loader=lambda: sd, device=dev)
    runner.trace()
    g = torch.Generator(device="cpu").manual_seed(5)
    prompts = torch.randint(1, 1024, (batch, 50), generator=g).to(dev)
    toks.append(runner.generate(prompts, max_new_tokens=new)[:, -new:].cpu())
    torch.cuda.synchronize()
    caches.append([(k.float().cpu().clone(), v.float().cpu().clone()) for k, v in
                   (runner.model.kv_cache.layer(i) for i in range(2))])
print("tokens differ at", (toks[0] != toks[1]).nonzero().tolist())
for li in range(2):
    for name, idx in (("K", 0), ("V", 1)):
        a, b = caches[0][li][idx], caches[1][li][idx]          # [B, Hkv, S, D]
        d = (a[:, :, :50 + new - 1] - b[:, :, :50 + new - 1]).abs()
        per = d.amax(dim=(1, 3))                                # [B, S]
        nz = (per > 0).nonzero()
        print(f"layer {li} {name}: max diff {d.max().item():.4f}; differing (row, pos) count {nz.shape[0]}; first {nz[:12].tolist()}")
