"""NxDPPModel over NCCL point-to-point on real GPUs (BASELINE config #3 in miniature).

PP = 2 on two GPUs: loss and gradients of the 1F1B and the interleaved schedule against the un-partitioned model on the
same GPU with the same kernels; TP2 x PP2 + sequence parallelism + activation checkpointing on four GPUs: the loss must
fall and be identical on every rank.  Skipped unless enough CUDA devices are visible.  Reference behaviour:
``src/neuronx_distributed/pipeline/model.py:773-830,1382-1397``, ``pipeline/comm.py:38-92``."""
from __future__ import annotations

import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _cfg(layers=4, **kw):
    from neuronx_distributed_llama3_2_b200.models import tiny_llama

    return tiny_llama(num_hidden_layers=layers, hidden_size=512, intermediate_size=1024, num_attention_heads=8,
                      num_key_value_heads=2, head_dim=64, vocab_size=1024, max_position_embeddings=1024,
                      torch_dtype=torch.bfloat16, **kw)


def _worker(rank, world, port, case, arg):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    try:
        globals()[case](rank, world, arg)
        torch.cuda.synchronize()
        dist.barrier()
    finally:
        ps.destroy_model_parallel()
        dist.destroy_process_group()


def _spawn(case, world, arg=None):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    from conftest import spawn_gpu_workers

    spawn_gpu_workers(_worker, (world, _free_port(), case, arg), world, deadline_s=240.0)


def case_pp2(rank, world, vp):
    from neuronx_distributed_llama3_2_b200 import trainer as nxd
    from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM
    from neuronx_distributed_llama3_2_b200.models.llama import LlamaDecoderLayer
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    dev = torch.device("cuda", rank)
    nxd_cfg = nxd.neuronx_distributed_config(
        tensor_parallel_size=1, pipeline_parallel_size=2,
        pipeline_config={"transformer_layer_cls": LlamaDecoderLayer, "num_microbatches": 4, "auto_partition": True,
                         "input_names": ["input_ids", "labels"], "virtual_pipeline_size": vp},
        optimizer_config={"zero_one_enabled": True, "grad_clipping": True, "max_grad_norm": 1.0})
    assert ps.get_pipeline_model_parallel_size() == 2
    cfg = _cfg(sequence_parallel_enabled=False)

    def build():
        torch.manual_seed(0)
        return LlamaForCausalLM(cfg, device=dev)

    ids = torch.randint(0, cfg.vocab_size, (8, 256), generator=torch.Generator().manual_seed(3)).to(dev)
    ref = build()                                       # oracle: the whole model on this GPU, same kernels
    ref_loss = torch.stack([ref(c, labels=c) for c in torch.chunk(ids, 4, dim=0)]).mean()
    ref_loss.backward()
    ref_grads = {n: p.grad.float().clone() for n, p in ref.named_parameters()}
    del ref

    model = nxd.initialize_parallel_model(nxd_cfg, build, device=dev)
    pp = model.module
    assert pp.num_stages == 2 * vp
    loss = model.run_train(input_ids=ids, labels=ids)
    if rank == world - 1:
        assert abs(loss.item() - ref_loss.item()) < 2e-3 * abs(ref_loss.item()), (loss.item(), ref_loss.item())
    else:
        assert loss is None
    for n, p in model.named_parameters():
        rn = "model.embed_tokens.weight" if n == "lm_head.weight" else n
        assert p.grad is not None and p.grad.is_cuda, n
        g, w = p.grad.float(), ref_grads[rn]
        rel = float((g - w).norm() / (w.norm() + 1e-12))
        assert rel < 3e-2, (n, rel)
    # optimizer steps through the trainer API: the loss falls, second step reuses the recorded tensor metadata
    opt = nxd.initialize_parallel_optimizer(nxd_cfg, torch.optim.AdamW, model.parameters(), lr=1e-3)
    losses = []
    for _ in range(4):
        opt.zero_grad()
        out = model.run_train(input_ids=ids, labels=ids)
        opt.step()
        if rank == world - 1:
            losses.append(float(out))
    if rank == world - 1:
        assert losses[-1] < losses[0], losses
    outs = model.run_eval(input_ids=ids, labels=ids)
    if rank == world - 1:
        assert len(outs) == 4


@pytest.mark.parametrize("vp", [1, 2])
def test_pp2_nccl_matches_unpartitioned_model(vp):
    _spawn("case_pp2", 2, vp)


def case_tp2_pp2(rank, world, _):
    """TP2 x PP2 + SP + full activation checkpointing + ZeRO-1: the fused AG->GEMM / GEMM->RS kernels run inside the
    pipeline stages, NCCL p2p carries [S/tp, B, H] activations between them."""
    import torch.distributed as dist

    from neuronx_distributed_llama3_2_b200 import trainer as nxd
    from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM
    from neuronx_distributed_llama3_2_b200.models.llama import LlamaDecoderLayer
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    dev = torch.device("cuda", rank)
    nxd_cfg = nxd.neuronx_distributed_config(
        tensor_parallel_size=2, pipeline_parallel_size=2, sequence_parallel=True,
        pipeline_config={"transformer_layer_cls": LlamaDecoderLayer, "num_microbatches": 2, "auto_partition": True,
                         "input_names": ["input_ids", "labels"], "broadcast_and_average_loss": True},
        optimizer_config={"zero_one_enabled": True, "grad_clipping": True, "max_grad_norm": 1.0},
        activation_checkpoint_config="full")
    assert (ps.get_tensor_model_parallel_size(), ps.get_pipeline_model_parallel_size()) == (2, 2)
    cfg = _cfg(sequence_parallel_enabled=True)
    torch.manual_seed(0)
    model = nxd.initialize_parallel_model(nxd_cfg, lambda: LlamaForCausalLM(cfg, device=dev), device=dev)
    opt = nxd.initialize_parallel_optimizer(nxd_cfg, torch.optim.AdamW, model.parameters(), lr=1e-3)
    ids = torch.randint(0, cfg.vocab_size, (4, 512), generator=torch.Generator().manual_seed(5)).to(dev)
    losses = []
    for _ in range(6):
        opt.zero_grad()
        losses.append(float(model.run_train(input_ids=ids, labels=ids)))
        opt.step()
    assert losses[-1] < losses[0] - 0.1, losses
    gathered = [None] * world
    dist.all_gather_object(gathered, losses)
    assert all(abs(g[-1] - losses[-1]) < 1e-4 for g in gathered), gathered


def test_tp2_pp2_sp_activation_checkpointing_nccl():
    _spawn("case_tp2_pp2", 4)
