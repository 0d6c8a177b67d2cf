"""pytest configuration: the ``gpu`` marker, CPU/gloo multi-process helper, per-test model-parallel reset."""
from __future__ import annotations

import os
import socket
import sys
import traceback

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, fn, args, errq):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    try:
        torch.manual_seed(1234)
        torch.set_num_threads(1)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        fn(rank, world, *args)
        dist.barrier()
    except Exception:
        errq.put((rank, traceback.format_exc()))
    finally:
        try:
            from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

            ps.destroy_model_parallel()
        except Exception:
            pass
        if dist.is_initialized():
            dist.destroy_process_group()


def run_distributed(fn, world: int, *args, timeout: float = 150.0):
    """Run ``fn(rank, world, *args)`` in ``world`` gloo processes on this host; re-raise the first failure."""
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    errq = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, fn, args, errq)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout)
    hung = [p for p in procs if p.is_alive()]
    for p in hung:
        p.kill()
    errs = []
    while not errq.empty():
        errs.append(errq.get())
    if errs:
        raise AssertionError("\n".join(f"[rank {r}]\n{tb}" for r, tb in errs))
    assert not hung, "distributed test timed out"
    for p in procs:
        assert p.exitcode == 0, f"worker exited with {p.exitcode}"


@pytest.fixture(autouse=True)
def _reset_parallel_state():
    yield
    try:
        from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

        if ps.model_parallel_is_initialized():
            ps.destroy_model_parallel()
    except Exception:
        pass


def spawn_gpu_workers(fn, args, nprocs: int, deadline_s: float = 150.0) -> None:
    """``torch.multiprocessing.spawn`` with a hard deadline: a rank that spins on a flag of a dead peer would otherwise
    hold the GPUs until the outer job limit.  On expiry the workers (exact PIDs) are killed and the test fails."""
    import os
    import signal
    import time

    import torch.multiprocessing as mp

    ctx = mp.spawn(fn, args=args, nprocs=nprocs, join=False)
    t0 = time.time()
    try:
        while not ctx.join(timeout=2.0):
            if time.time() - t0 > deadline_s:
                raise TimeoutError(f"GPU workers still running after {deadline_s:.0f} s")
    except BaseException:
        for proc in ctx.processes:
            if proc.is_alive():
                os.kill(proc.pid, signal.SIGKILL)
        raise
