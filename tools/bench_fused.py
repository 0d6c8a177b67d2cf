#!/usr/bin/env python
"""Fused AG->GEMM (and later GEMM->RS) against NCCL + the same tcgen05 GEMM, at Llama-3.2-3B tensor-parallel shapes.
Launch: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/bench_fused.py
Device-timed (CUDA events), max over ranks."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    from neuronx_distributed_llama3_2_b200.ops.fused_comm import fused_ag_gemm
    from neuronx_distributed_llama3_2_b200.ops.gemm import gemm
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    ps.initialize_model_parallel(tensor_model_parallel_size=world)
    ctx = ps.get_symm_mem_context("tp")
    S, H, I = int(os.environ.get("NXD_BENCH_S", 8192)), 3072, 8192
    heads = 24 if world <= 8 else 32
    R = S // world
    cases = [("qkv fwd", H, (24 + 16) * 128 // world), ("gate_up fwd", H, 2 * I // world)]

    def timed(fn, iters=20, warmup=5):
        for _ in range(warmup):
            fn()
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    for name, K, N in cases:
        x = torch.randn(R, K, device="cuda").bfloat16()
        w = torch.randn(N, K, device="cuda").bfloat16()
        total = torch.empty(world * R, K, device="cuda", dtype=torch.bfloat16)

        def nccl():
            dist.all_gather_into_tensor(total, x)
            return gemm(total, w, "nt")

        def gemm_only():
            return gemm(total, w, "nt")

        def ag_only():
            dist.all_gather_into_tensor(total, x)

        def fused():
            return fused_ag_gemm(ctx, x, w, "nt", persistent_total=False)

        res = {"case": name, "tp": world, "R": R, "K": K, "N": N, "nccl_ag_plus_gemm_ms": timed(nccl),
               "gemm_only_ms": timed(gemm_only), "nccl_ag_only_ms": timed(ag_only), "fused_ms": timed(fused)}
        bytes_in = (world - 1) * R * K * 2
        flops = 2.0 * world * R * K * N
        res["roofline_ms"] = max(flops / 1373.1e12, bytes_in / 770e9) * 1e3
        res["fused_frac_of_roofline"] = res["roofline_ms"] / res["fused_ms"]
        res["speedup_vs_nccl"] = res["nccl_ag_plus_gemm_ms"] / res["fused_ms"]
        if rank == 0:
            print(json.dumps(res), flush=True)
    from neuronx_distributed_llama3_2_b200.ops.fused_comm import fused_gemm_rs

    for name, K, N in [("o_proj fwd", 24 * 128 // world, H), ("down fwd", I // world, H)]:
        x = torch.randn(S, K, device="cuda").bfloat16()
        w = torch.randn(N, K, device="cuda").bfloat16()
        out = torch.empty(R, N, device="cuda", dtype=torch.bfloat16)

        def nccl():
            y = gemm(x, w, "nt")
            dist.reduce_scatter_tensor(out, y)

        def gemm_only():
            return gemm(x, w, "nt")

        yfull = gemm(x, w, "nt")

        def rs_only():
            dist.reduce_scatter_tensor(out, yfull)

        def fused():
            return fused_gemm_rs(ctx, x, w, "nt")

        res = {"case": name, "tp": world, "R": R, "K": K, "N": N, "gemm_plus_nccl_rs_ms": timed(nccl),
               "gemm_only_ms": timed(gemm_only), "nccl_rs_only_ms": timed(rs_only), "fused_ms": timed(fused)}
        bytes_out = (world - 1) * R * N * 2
        flops = 2.0 * S * K * N
        res["roofline_ms"] = max(flops / 1373.1e12, bytes_out / 770e9) * 1e3
        res["fused_frac_of_roofline"] = res["roofline_ms"] / res["fused_ms"]
        res["speedup_vs_nccl"] = res["gemm_plus_nccl_rs_ms"] / res["fused_ms"]
        if rank == 0:
            print(json.dumps(res), flush=True)
    if os.environ.get("NXD_FUSED_TRACE", "0") == "1" and rank == 0:
        from neuronx_distributed_llama3_2_b200.ops.fused_comm import dump_traces

        for what, stamps in dump_traces(400)[::20]:
            print("TRACE", what, stamps, flush=True)
    ps.destroy_model_parallel()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
