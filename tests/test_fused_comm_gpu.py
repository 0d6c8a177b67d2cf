"""Fused compute+collective kernels across real GPUs (one process per GPU over NCCL + CUDA-IPC symmetric memory).
Skipped unless at least two CUDA devices are visible."""
from __future__ import annotations

import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu



def _backend():
    from neuronx_distributed_llama3_2_b200.ops import backend

    return backend

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
    from conftest import spawn_gpu_workers

    spawn_gpu_workers(_worker, (world, _free_port(), case), world)


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
        os.environ["NXD_AG_BLOCK_N"] = ("128", "256")[it % 2]      # both tile widths (the default is a wave heuristic)
        res = fused_ag_gemm(ctx, x, w, layout)
        os.environ.pop("NXD_AG_BLOCK_N")
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


def case_gemm_ar(rank, world):
    import torch.distributed as dist

    from neuronx_distributed_llama3_2_b200.ops.fused_comm import fused_gemm_ar, oneshot_all_reduce, skinny_gemm
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    ctx = ps.get_symm_mem_context("tp")
    # token generation shapes (skinny GEMM -> one-shot all-reduce in one kernel), repeated to exercise both parities
    for it, (B, K, N) in enumerate([(4, 1024, 3072), (1, 512, 2048), (3, 1024, 3080), (4, 1024, 3072),
                                    (8, 384, 3072)]):
        g = torch.Generator(device="cuda").manual_seed(31 * it + rank)
        x = torch.randn(B, K, device="cuda", generator=g).bfloat16()
        w = (torch.randn(N, K, device="cuda", generator=g) * 0.05).bfloat16()
        out = fused_gemm_ar(ctx, x, w, "nt")
        assert out is not None and out.shape == (B, N)
        ref = x.float() @ w.float().t()
        dist.all_reduce(ref)
        assert _rel(out, ref) < 4e-3, (it, _rel(out, ref))
        gathered = [torch.empty_like(out) for _ in range(world)]
        dist.all_gather(gathered, out)
        assert all(torch.equal(gathered[0], t) for t in gathered), "ranks disagree bitwise"
    # with the residual add fused into the epilogue
    x = torch.randn(4, 1024, device="cuda").bfloat16()
    w = (torch.randn(3072, 1024, device="cuda") * 0.05).bfloat16()
    res = torch.randn(4, 3072, device="cuda").bfloat16()
    dist.broadcast(res, 0)
    out = skinny_gemm(x, w, residual=res, ctx=ctx)
    ref = x.float() @ w.float().t()
    dist.all_reduce(ref)
    assert _rel(out, ref + res.float()) < 4e-3
    # short-prefill shape: tcgen05 GEMM + one-shot all-reduce
    x = torch.randn(128, 1024, device="cuda").bfloat16()
    out = fused_gemm_ar(ctx, x, w, "nt")
    ref = x.float() @ w.float().t()
    dist.all_reduce(ref)
    assert out is not None and _rel(out, ref) < 8e-3
    # stand-alone one-shot all-reduce, several sizes
    for n in (8, 4096, 24576, 1 << 19, 1000 * 8):
        t = torch.randn(n, device="cuda").bfloat16()
        got = oneshot_all_reduce(ctx, t)
        ref = t.float()
        dist.all_reduce(ref)
        assert got is not None and _rel(got, ref) < 4e-3, n


def case_sp_embedding(rank, world):
    import torch.distributed as dist

    from neuronx_distributed_llama3_2_b200.parallel.layers import ParallelEmbedding

    V, H, B, S = 1024, 512, 2, 256
    emb = ParallelEmbedding(V, H, dtype=torch.bfloat16, device=torch.device("cuda"), sequence_parallel_enabled=True)
    full = torch.empty(V, H, device="cuda", dtype=torch.bfloat16)
    dist.all_gather_into_tensor(full, emb.weight.data.contiguous())
    for it in range(4):
        g = torch.Generator(device="cpu").manual_seed(50 + it)
        ids = torch.randint(0, V, (B, S), generator=g).cuda()
        _backend().set_backends(fused_comm=True)
        out = emb(ids)                                   # [S/tp, B, H] through the push kernel
        ref = full[ids].transpose(0, 1)[rank * (S // world):(rank + 1) * (S // world)]
        assert out.shape == ref.shape and torch.equal(out, ref), it
        dy = torch.randn_like(out)
        emb.weight.grad = None
        out.backward(dy)
        g_push = emb.weight.grad.clone()
        _backend().set_backends(fused_comm=False)
        emb.weight.grad = None
        out2 = emb(ids)                                  # NCCL reduce-scatter path
        assert torch.equal(out2, ref)
        out2.backward(dy)
        assert _rel(g_push, emb.weight.grad) < 1e-3
    _backend().set_backends(fused_comm=True)


def _worker_dp(rank, world, port, case):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    ps.initialize_model_parallel(tensor_model_parallel_size=1)      # pure data parallel: dp = world
    try:
        globals()[case](rank, world)
        torch.cuda.synchronize()
        dist.barrier()
    finally:
        ps.destroy_model_parallel()
        dist.destroy_process_group()


def case_zero1_fused(rank, world):
    """Fused ZeRO-1 step (pull reduce-scatter + AdamW with pushed all-gather) against the NCCL path."""
    import torch.distributed as dist

    from neuronx_distributed_llama3_2_b200.optimizer import NeuronZero1Optimizer

    def run(fused: bool):
        os.environ["NXD_FUSED_ZERO1"] = "1" if fused else "0"
        torch.manual_seed(0)
        model = torch.nn.Sequential(torch.nn.Linear(512, 1024, bias=False), torch.nn.GELU(),
                                    torch.nn.Linear(1024, 384, bias=True)).cuda().bfloat16()
        opt = NeuronZero1Optimizer(list(model.parameters()), torch.optim.AdamW, lr=1e-2, betas=(0.9, 0.95),
                                   weight_decay=0.1, grad_clipping=True, max_norm=1.0)
        assert (opt.buckets[0].symm is not None) == fused
        norms = []
        for step in range(4):
            g = torch.Generator(device="cuda").manual_seed(100 * step + rank)      # different data per DP rank
            x = torch.randn(64, 512, device="cuda", generator=g).bfloat16()
            opt.zero_grad()
            model(x).float().pow(2).mean().backward()
            opt.step()
            norms.append(float(opt.grad_norm))
        flat = torch.cat([p.detach().float().reshape(-1) for p in model.parameters()])
        return flat, norms

    ref, ref_norms = run(False)
    got, got_norms = run(True)
    os.environ["NXD_FUSED_ZERO1"] = "1"
    # every rank holds the same parameters after the pushed all-gather
    both = [torch.empty_like(got) for _ in range(world)]
    dist.all_gather(both, got)
    assert all(torch.equal(both[0], t) for t in both)
    for a, b in zip(got_norms, ref_norms):
        assert abs(a - b) < 2e-2 * max(1.0, b), (got_norms, ref_norms)
    # fp32 summation across ranks vs NCCL's bf16 ring: agreement to bf16 resolution of the update
    assert _rel(got, ref) < 5e-3, _rel(got, ref)


def _worker_ep(rank, world, port, case):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    ps.initialize_model_parallel(tensor_model_parallel_size=1, expert_model_parallel_size=world)
    try:
        globals()[case](rank, world)
        torch.cuda.synchronize()
        dist.barrier()
    finally:
        ps.destroy_model_parallel()
        dist.destroy_process_group()


def case_moe_fused(rank, world):
    """Expert-parallel dispatch + GEMM and GEMM + combine kernels against the NCCL all-to-all + per-expert GEMM path."""
    from neuronx_distributed_llama3_2_b200.modules.moe.experts import Experts

    E, C, H, I = 4, 256, 256, 512
    torch.manual_seed(5)
    experts = Experts(E, H, I, glu=True, activation_fn=torch.nn.functional.silu, dtype=torch.bfloat16,
                      device=torch.device("cuda"))
    with torch.no_grad():
        for p in experts.parameters():
            p.mul_(0.5)
    res = {}
    for it in range(3):                                   # three rounds: both buffer parities are reused
        g = torch.Generator(device="cuda").manual_seed(70 + 10 * it + rank)
        x = torch.randn(E, C, H, device="cuda", generator=g).bfloat16()
        dy = torch.randn(E, C, H, device="cuda", generator=g).bfloat16()
        for fused in ("0", "1"):
            _backend().set_backends(fused_moe=(fused == "1"))
            xr = x.clone().requires_grad_()
            for p in experts.parameters():
                p.grad = None
            out = experts(xr)
            out.backward(dy)
            res[fused] = (out.detach(), xr.grad.clone(), [p.grad.clone() for p in experts.parameters()])
        o0, gx0, gw0 = res["0"]
        o1, gx1, gw1 = res["1"]
        assert _rel(o1, o0) < 1e-2, (it, _rel(o1, o0))
        assert _rel(gx1, gx0) < 1e-2, (it, _rel(gx1, gx0))
        for a, b in zip(gw1, gw0):
            assert _rel(a, b) < 1e-2, (it, _rel(a, b))
    _backend().set_backends(fused_moe=True)


def case_flag_protocol_stress(rank, world):
    """Back-to-back calls of the four symmetric-memory kernels with changing payloads and skewed rank arrival (a
    busy-wait kernel delays a different rank every iteration), each checked against NCCL.  A flag that is consumed one
    call early / late, a staging buffer reused before its readers are done, or a missing system fence shows up as a
    stale payload (the generator seed changes every call) or as a hang (bounded by the spawn deadline)."""
    import torch.distributed as dist

    from neuronx_distributed_llama3_2_b200.ops.fused_comm import (fused_ag_gemm, fused_gemm_ar, fused_gemm_rs,
                                                                    oneshot_all_reduce)
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    ctx = ps.get_symm_mem_context("tp")
    iters = int(os.environ.get("NXD_STRESS_ITERS", "200"))
    R, K, N = 256, 512, 768
    w_nt = (torch.randn(N, K, device="cuda", generator=torch.Generator(device="cuda").manual_seed(3)) * 0.1).bfloat16()
    for it in range(iters):
        g = torch.Generator(device="cuda").manual_seed(7919 * it + rank)
        if it % world == rank:                                   # this rank arrives late (~0.2 ms at 2 GHz)
            torch.cuda._sleep(400_000 * (1 + it % 3))
        op = it % 4
        if op == 0:
            x = torch.randn(R, K, device="cuda", generator=g).bfloat16()
            y, x_total = fused_ag_gemm(ctx, x, w_nt, "nt")
            ref_total = torch.empty(world * R, K, device="cuda", dtype=torch.bfloat16)
            dist.all_gather_into_tensor(ref_total, x)
            assert torch.equal(x_total, ref_total), f"ag iteration {it}: stale or torn gather"
            assert _rel(y, ref_total.float() @ w_nt.float().t()) < 6e-3, it
        elif op == 1:
            x = torch.randn(world * R, K, device="cuda", generator=g).bfloat16()
            out = fused_gemm_rs(ctx, x, w_nt, "nt")
            ref = x.float() @ w_nt.float().t()
            dist.all_reduce(ref)
            assert _rel(out, ref[rank * R:(rank + 1) * R]) < 8e-3, f"rs iteration {it}"
        elif op == 2:
            x = torch.randn(1 + it % 8, K, device="cuda", generator=g).bfloat16()
            out = fused_gemm_ar(ctx, x, w_nt, "nt")
            ref = x.float() @ w_nt.float().t()
            dist.all_reduce(ref)
            assert _rel(out, ref) < 4e-3, f"gemm_ar iteration {it}"
        else:
            t = torch.randn(8 * (1 + (37 * it) % 4096), device="cuda", generator=g).bfloat16()
            got = oneshot_all_reduce(ctx, t)
            ref = t.float()
            dist.all_reduce(ref)
            assert _rel(got, ref) < 4e-3, f"oneshot iteration {it}"


@pytest.mark.skipif(os.environ.get("NXD_STRESS", "0") != "1", reason="opt-in: NXD_STRESS=1 (see tools/run_sanitizers.sh)")
def test_flag_protocol_stress():
    world = int(os.environ.get("NXD_STRESS_WORLD", "2"))
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    from conftest import spawn_gpu_workers

    spawn_gpu_workers(_worker, (world, _free_port(), "case_flag_protocol_stress"), world, deadline_s=300.0)


# every protocol at world 2, 4 and 8 (7 peers, all copy CTAs): skipped where the box has fewer GPUs
_WORLDS = [2, 4, 8]


@pytest.mark.parametrize("world", _WORLDS)
def test_fused_ag_gemm(world):
    _spawn("case_ag_gemm", world)


@pytest.mark.parametrize("world", _WORLDS)
def test_fused_gemm_rs(world):
    _spawn("case_gemm_rs", world)


@pytest.mark.parametrize("world", _WORLDS)
def test_fused_gemm_ar(world):
    _spawn("case_gemm_ar", world)


@pytest.mark.parametrize("world", _WORLDS)
def test_sp_embedding_push(world):
    _spawn("case_sp_embedding", world)


def test_zero1_fused_step_dp2():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from conftest import spawn_gpu_workers

    spawn_gpu_workers(_worker_dp, (2, _free_port(), "case_zero1_fused"), 2)


def test_moe_fused_dispatch_combine_ep2():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from conftest import spawn_gpu_workers

    spawn_gpu_workers(_worker_ep, (2, _free_port(), "case_moe_fused"), 2, deadline_s=90.0)
