#!/usr/bin/env python
"""NVLink push-bandwidth sweep (csrc/p2p_probe.cu): every rank pushes simultaneously, device-timed, max over ranks.
Launch: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/bench_p2p.py"""
import ctypes as _c
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
    from neuronx_distributed_llama3_2_b200.ops import _lib
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    ps.initialize_model_parallel(tensor_model_parallel_size=world)
    ctx = ps.get_symm_mem_context("tp")
    lib = _lib.lib()
    lib.nxd_p2p_push_probe.argtypes = [_c.c_int, _c.c_void_p, _c.c_void_p, _c.c_int, _c.c_size_t, _c.c_int, _c.c_int,
                                       _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_size_t, _c.c_void_p]
    shard = 1024 * 3072 * 2                       # one tp8 activation shard (6.3 MB)
    total = shard * 8
    local, table = ctx.region("p2p_probe", total * 2)
    peers = [int(v) for v in table.tolist()]
    src = torch.randn(total // 2, device="cuda").bfloat16()
    nccl_out = torch.empty(world * shard // 2, device="cuda", dtype=torch.bfloat16)

    def timed(fn, iters=10, warmup=3):
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

    def push(mode, n_dst, nbytes, ctas, chunk=16384, bufs=12, lag=4, sig_mode=0, sig_every=4):
        # destination d = rank + 1 + d (ring order); every writer owns region [rank * nbytes, ...) of the peer buffer
        dsts = (_c.c_void_p * n_dst)(*[peers[(rank + 1 + d) % world] + (rank * shard if n_dst > 1 else 0) for d in range(n_dst)])

        def fn():
            _lib.check(lib.nxd_p2p_push_probe(mode, src.data_ptr(), dsts, n_dst, nbytes, ctas, chunk, bufs, lag, sig_mode,
                                              sig_every, total + rank * 64, _lib.stream_ptr()), "probe")
        return fn

    def report(tag, ms, out_bytes):
        if rank == 0:
            print(json.dumps({"case": tag, "ms": round(ms, 4), "out_GBps": round(out_bytes / ms / 1e6, 1)}), flush=True)

    ag = timed(lambda: dist.all_gather_into_tensor(nccl_out, src[: shard // 2]))
    report(f"nccl all_gather shard={shard}", ag, (world - 1) * shard)
    nd_all = world - 1
    quick = os.environ.get("NXD_P2P_FULL", "0") != "1"
    for n_dst, nbytes in ((1, nd_all * shard), (nd_all, shard)) if world > 2 else ((1, shard * 4),):
        if not quick:
            for mode in (0, 2, 1):
                for ctas in (8, 16, 24, 32, 48, 64, 96, 132):
                    ms = timed(push(mode, n_dst, nbytes, ctas))
                    report(f"mode{mode} n_dst={n_dst} ctas={ctas}", ms, n_dst * nbytes)
            for chunk, bufs, lag in ((4096, 32, 8), (8192, 24, 8), (32768, 6, 2), (16384, 12, 1), (16384, 4, 1), (2048, 32, 8)):
                for ctas in (24, 64):
                    ms = timed(push(0, n_dst, nbytes, ctas, chunk, bufs, lag))
                    report(f"mode0 n_dst={n_dst} ctas={ctas} chunk={chunk} bufs={bufs} lag={lag}", ms, n_dst * nbytes)
        # cost of signalling "these bytes have landed" to the peers while the stream keeps running
        for sig_mode in (0, 1, 2, 3, 4):
            for sig_every in (1, 4, 16):
                for ctas in (24, 48):
                    ms = timed(push(0, n_dst, nbytes, ctas, sig_mode=sig_mode, sig_every=sig_every))
                    report(f"mode0 n_dst={n_dst} ctas={ctas} sig_mode={sig_mode} sig_every={sig_every}", ms, n_dst * nbytes)
    ps.destroy_model_parallel()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
