"""Fused compute+collective kernels across real GPUs (one process per GPU over NCCL + CUDA-IPC symmetric memory).
Skipped unless at least two CUDA devices are visible."""
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


def _worker(rank, world, port, case):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    ps.initialize_model_parallel(tensor_model_parallel_size=world)
    try:
        globals()[case](rank, world)
        torch.cuda.synchronize()
        dist.barrier()
    finally:
        ps.destroy_model_parallel()
        dist.destroy_process_group()


def _spawn(case, world=2):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp

    mp.spawn(_worker, args=(world, _free_port(), case), nprocs=world, join=True)


def _rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


def case_ag_gemm(rank, world):
    import torch.distributed as dist

    from neuronx_distributed_llama3_2_b200.ops.fused_comm import fused_ag_gemm
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    ctx = ps.get_symm_mem_context("tp")
    assert ctx is not None and ctx.world == world
    for it, (R, K, N, layout) in enumerate([(512, 1024, 768, "nt"), (512, 1024, 768, "nt"), (512, 1024, 768, "nn"),
                                            (512, 1024, 2048, "nt"), (512, 1024, 768, "nt")]):
        g = torch.Generator(device="cuda").manual_seed(100 * it + rank)
        x = torch.randn(R, K, device="cuda", generator=g).bfloat16()
        gw = torch.Generator(device="cuda").manual_seed(7 + it)
        w = torch.randn((N, K) if layout == "nt" else (K, N), device="cuda", generator=gw).bfloat16()
        res = fused_ag_gemm(ctx, x, w, layout)
        assert res is not None
        y, x_total = res
        ref_total = torch.empty(world * R, K, device="cuda", dtype=torch.bfloat16)
        dist.all_gather_into_tensor(ref_total, x)
        assert torch.equal(x_total, ref_total), f"iteration {it}: gathered activations differ"
        ref = ref_total.float() @ (w.float().t() if layout == "nt" else w.float())
        assert _rel(y, ref) < 6e-3, (it, _rel(y, ref))


def case_gemm_rs(rank, world):
    import torch.distributed as dist

    from neuronx_distributed_llama3_2_b200.ops.fused_comm import fused_gemm_rs
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    ctx = ps.get_symm_mem_context("tp")
    for it, (R, K, N, layout) in enumerate([(512, 1024, 768, "nt"), (512, 1024, 768, "nt"), (512, 1024, 768, "nt"),
                                            (512, 512, 1024, "nn"), (256, 2048, 3072, "nt"), (512, 1024, 768, "nt")]):
        g = torch.Generator(device="cuda").manual_seed(1000 * it + rank)
        x = torch.randn(world * R, K, device="cuda", generator=g).bfloat16()
        w = torch.randn((N, K) if layout == "nt" else (K, N), device="cuda", generator=g).bfloat16()
        out = fused_gemm_rs(ctx, x, w, layout)
        assert out is not None and out.shape == (R, N)
        partial = x.float() @ (w.float().t() if layout == "nt" else w.float())
        dist.all_reduce(partial)
        ref = partial[rank * R:(rank + 1) * R]
        assert _rel(out, ref) < 8e-3, (it, _rel(out, ref))


def test_fused_ag_gemm_tp2():
    _spawn("case_ag_gemm", 2)


def test_fused_gemm_rs_tp2():
    _spawn("case_gemm_rs", 2)
