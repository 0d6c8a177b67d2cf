"""Small helpers shared by ``__graft_entry__.smoke`` and the examples."""
from __future__ import annotations

import torch


def smoke_train_step(device: str = "cuda:0", steps: int = 2) -> float:
    """Tiny Llama-3.2-style model (4 layers, hidden 512, GQA 8/2 heads of 64) through the public trainer API:
    forward + backward + ZeRO-1 AdamW step, all hot ops on the sm_100a kernels."""
    from . import trainer as nxd
    from .models import LlamaConfig, LlamaForCausalLM

    dev = torch.device(device)
    torch.cuda.set_device(dev)
    cfg = LlamaConfig(vocab_size=4096, hidden_size=512, intermediate_size=1536, num_hidden_layers=4,
                      num_attention_heads=8, num_key_value_heads=2, head_dim=64, max_position_embeddings=2048,
                      sequence_parallel_enabled=False)
    nxd_cfg = nxd.neuronx_distributed_config(optimizer_config={"zero_one_enabled": True, "grad_clipping": True,
                                                               "max_grad_norm": 1.0},
                                             mixed_precision_config={"use_master_weights": True,
                                                                     "use_fp32_grad_acc": False,
                                                                     "use_master_weights_in_ckpt": False})
    torch.manual_seed(0)
    model = nxd.initialize_parallel_model(nxd_cfg, lambda: LlamaForCausalLM(cfg, device=dev), device=dev)
    opt = nxd.initialize_parallel_optimizer(nxd_cfg, torch.optim.AdamW, model.parameters(), lr=1e-3)
    ids = torch.randint(0, cfg.vocab_size, (2, 512), device=dev)
    loss = None
    first = None
    for _ in range(steps):
        opt.zero_grad()
        loss = model.run_train(ids, labels=ids)
        opt.step()
        first = first if first is not None else float(loss)
    final = float(loss)
    assert final == final and abs(final) < 50, f"bad loss {final}"
    return final
