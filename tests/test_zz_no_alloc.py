"""The allocation-free RMSNorm backward entry point (``nxd_rmsnorm_bwd_static``).

It exists for callers that cannot hand the library a scratch buffer -- a process whose device memory is owned by
another allocator, a loop that must not hit cudaMalloc's synchronisation, a plug-in that may not allocate at all.
The test drives it straight through the C ABI (ctypes, raw device pointers), the way such a caller would, and checks
it against the ordinary entry point that takes a caller-provided scratch buffer.  Two calls in a row with different
data make sure nothing stale in the shared built-in scratch leaks into a later result.

This file is named so that pytest collects it last: a ``-x`` run has already exercised every other test when it
gets here.  It passes on its own in a fresh process as well as after the rest of the suite.
"""
from __future__ import annotations

import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


def _ptr(t):
    return C.c_void_p(0 if t is None else t.data_ptr())


def _run_pair(lib, rows, H, seed, with_residual):
    g = torch.Generator(device="cuda").manual_seed(seed)
    dy = torch.randn(rows, H, device="cuda", generator=g).bfloat16()
    x = torch.randn(rows, H, device="cuda", generator=g).bfloat16()
    w = (1 + 0.1 * torch.randn(H, device="cuda", generator=g)).bfloat16()
    dres = torch.randn(rows, H, device="cuda", generator=g).bfloat16() if with_residual else None
    rstd = torch.rsqrt(x.float().pow(2).mean(-1) + 1e-5).contiguous()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    # existing entry point: the caller owns the scratch
    P = lib.nxd_rmsnorm_bwd_num_partials(C.c_int(rows))
    scratch = torch.empty(P * H, dtype=torch.float32, device="cuda")
    dx_ref = torch.empty_like(x)
    dw_ref = torch.empty(H, dtype=torch.float32, device="cuda")
    rc = lib.nxd_rmsnorm_bwd(_ptr(dy), _ptr(x), _ptr(w), _ptr(rstd), _ptr(dres), _ptr(dx_ref), _ptr(None), _ptr(dw_ref),
                             C.c_int(0), _ptr(scratch), C.c_int(rows), C.c_int(H), stream)
    assert rc == 0, lib.nxd_last_error()

    # new entry point: inputs and outputs only
    dx = torch.empty_like(x)
    dw = torch.empty(H, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    rc = lib.nxd_rmsnorm_bwd_static(_ptr(dy), _ptr(x), _ptr(w), _ptr(rstd), _ptr(dres), _ptr(dx), _ptr(None), _ptr(dw),
                                    C.c_int(0), C.c_int(rows), C.c_int(H), stream)
    assert rc == 0, lib.nxd_last_error()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    assert torch.equal(dx, dx_ref)                 # dx does not depend on the partial-sum layout at all
    assert _rel(dw, dw_ref) < 8e-3                 # same tolerance as test_rmsnorm_fwd_bwd
    # and against plain fp32 math
    xf, wf = x.float(), w.float()
    xh = xf * rstd[:, None]
    assert _rel(dw, (dy.float() * xh).sum(0)) < 8e-3


def test_rmsnorm_bwd_static_matches_scratch_version_twice():
    from neuronx_distributed_llama3_2_b200.ops import _lib

    assert _lib.available()
    lib = _lib.lib()
    lib.nxd_last_error.restype = C.c_char_p
    assert lib.nxd_rmsnorm_bwd_static_max_h() == 8192
    _run_pair(lib, rows=4096, H=3072, seed=1, with_residual=False)
    _run_pair(lib, rows=777, H=8192, seed=2, with_residual=True)     # second call: larger H, fewer partial rows
    _run_pair(lib, rows=5, H=256, seed=3, with_residual=False)       # and a tiny one over the same stale scratch


def test_rmsnorm_bwd_static_rejects_oversize():
    from neuronx_distributed_llama3_2_b200.ops import _lib

    lib = _lib.lib()
    lib.nxd_last_error.restype = C.c_char_p
    H = 16384
    t = torch.zeros(2, H, device="cuda", dtype=torch.bfloat16)
    w = torch.ones(H, device="cuda", dtype=torch.bfloat16)
    rstd = torch.ones(2, device="cuda")
    dw = torch.zeros(H, device="cuda")
    rc = lib.nxd_rmsnorm_bwd_static(_ptr(t), _ptr(t), _ptr(w), _ptr(rstd), _ptr(None), _ptr(t.clone()), _ptr(None),
                                    _ptr(dw), C.c_int(0), C.c_int(2), C.c_int(H),
                                    C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc != 0
    assert b"exceeds the built-in scratch limit" in lib.nxd_last_error()
