"""Convergence-curve parity on the GPU (reference ``test/integration/combinatorial_tests/run.sh:142-160`` with
``common/compare_gpu_trn1_metrics.py:43-50,71-73``: 50 step losses, at least 95 % of them within rtol 5 % of a stored GPU
run).  Here the stored run is the LIBRARY arm of the same model (cuBLAS GEMMs + SDPA attention) trained in the same
process from the same weights on the same batches; the candidate is the native arm (tcgen05 GEMMs + our attention)."""
from __future__ import annotations

import pytest
import torch

pytestmark = pytest.mark.gpu

STEPS = 50


def _run(arm):
    from neuronx_distributed_llama3_2_b200 import trainer as nxd
    from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM, tiny_llama
    from neuronx_distributed_llama3_2_b200.ops import backend
    from neuronx_distributed_llama3_2_b200.parallel.random import model_parallel_manual_seed

    dev = torch.device("cuda", 0)
    cfg = tiny_llama(num_hidden_layers=4, hidden_size=1024, intermediate_size=2816, num_attention_heads=8,
                     num_key_value_heads=2, head_dim=128, vocab_size=8192, max_position_embeddings=2048,
                     torch_dtype=torch.bfloat16, sequence_parallel_enabled=False)
    with backend.use_backends(**arm):
        nxd_cfg = nxd.neuronx_distributed_config(
            optimizer_config={"zero_one_enabled": True, "grad_clipping": True, "max_grad_norm": 1.0})
        model_parallel_manual_seed(4321)
        torch.manual_seed(4321)
        model = nxd.initialize_parallel_model(nxd_cfg, lambda: LlamaForCausalLM(cfg, device=dev), device=dev)
        model.train()
        opt = nxd.initialize_parallel_optimizer(nxd_cfg, torch.optim.AdamW, model.parameters(), lr=3e-4,
                                                betas=(0.9, 0.95), weight_decay=0.1)
        g = torch.Generator().manual_seed(77)
        # a learnable, non-memorisable signal: i.i.d. tokens from a Zipf-like unigram distribution (the loss falls smoothly
        # from ln(V) towards the entropy of the distribution, as a real pre-training curve does in its first steps)
        probs = 1.0 / (torch.arange(cfg.vocab_size, dtype=torch.float64) + 10.0)
        probs /= probs.sum()
        losses = []
        for step in range(STEPS):
            ids = torch.multinomial(probs, 2 * 1024, replacement=True, generator=g).reshape(2, 1024)
            ids = ids.to(dev)
            for grp in opt.param_groups:                      # linear warm-up over 10 steps, as every real recipe has
                grp["lr"] = 3e-4 * min(1.0, (step + 1) / 10)
            opt.zero_grad()
            loss = model.run_train(ids, labels=ids)
            opt.step()
            losses.append(float(loss))
        counts = backend.fallback_counts()
    return losses, counts


def test_native_arm_follows_the_library_arm_for_50_steps():
    from neuronx_distributed_llama3_2_b200.ops import backend
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    backend.reset_fallback_counts()
    native, fb = _run(backend.NATIVE_ARM)
    ps.destroy_model_parallel()
    library, _ = _run(backend.LIBRARY_ARM)
    assert not fb, f"the native arm fell back to library kernels: {fb}"
    assert native[-1] < native[0] - 0.5, ("the model did not learn", native[0], native[-1])
    close = sum(1 for a, b in zip(native, library) if abs(a - b) <= 0.05 * abs(b))
    assert close >= 0.95 * STEPS, (close, list(zip(native, library)))
