"""Numerics of every hand-written sm_100a kernel against a plain fp32 PyTorch restatement of the same op.

Run on the B200 box:  python -m pytest tests -m gpu -x -q
"""
from __future__ import annotations

import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from neuronx_distributed_llama3_2_b200.ops import _lib

    assert _lib.available(), "libnxd_b200.so must load on a GPU box"
    return _lib


def _rel_err(a: torch.Tensor, b: torch.Tensor) -> float:
    a, b = a.float(), b.float()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


# ---------------------------------------------------------------------------------------------
# tcgen05 GEMM
# ---------------------------------------------------------------------------------------------

_SHAPES = [
    (256, 256, 64),
    (256, 128, 128),
    (512, 384, 512),
    (1024, 1024, 1024),
    (8192, 2048, 3072),
    (4096, 3072, 8192),
    (1000, 264, 328),     # ragged M/N/K (multiples of 8 only)
    (128, 5120, 3072),
]


@pytest.mark.parametrize("layout", ["nt", "nn", "tn"])
@pytest.mark.parametrize("shape", _SHAPES)
def test_gemm_layouts(layout, shape):
    from neuronx_distributed_llama3_2_b200.ops.gemm import gemm

    _lib()
    M, N, K = shape
    g = torch.Generator(device="cuda").manual_seed(M * 7 + N * 3 + K)
    if layout == "nt":
        a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
        b = torch.randn(N, K, device="cuda", generator=g).bfloat16()
        ref = a.float() @ b.float().t()
    elif layout == "nn":
        a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
        b = torch.randn(K, N, device="cuda", generator=g).bfloat16()
        ref = a.float() @ b.float()
    else:
        a = torch.randn(K, M, device="cuda", generator=g).bfloat16()
        b = torch.randn(K, N, device="cuda", generator=g).bfloat16()
        ref = a.float().t() @ b.float()
    for bn in (128, 256):
        out = gemm(a, b, layout, block_n=bn)
        torch.cuda.synchronize()
        assert _rel_err(out, ref) < 6e-3, (layout, shape, bn, _rel_err(out, ref))
    out32 = gemm(a, b, layout, out_dtype=torch.float32)
    assert _rel_err(out32, ref) < 1e-4
    acc = torch.ones(M, N, device="cuda", dtype=torch.float32)
    gemm(a, b, layout, out=acc, accumulate=True)
    assert _rel_err(acc, ref + 1.0) < 1e-4


def test_gemm_strided_operands():
    from neuronx_distributed_llama3_2_b200.ops.gemm import gemm

    _lib()
    big = torch.randn(512, 1024, device="cuda").bfloat16()
    a = big[:, 256:768]                        # row pitch 1024, K = 512
    w = torch.randn(384, 512, device="cuda").bfloat16()
    out_big = torch.zeros(512, 1024, device="cuda", dtype=torch.bfloat16)
    out = out_big[:, 128:512]
    gemm(a, w, "nt", out=out)
    ref = a.float() @ w.float().t()
    assert _rel_err(out, ref) < 6e-3
    assert out_big[:, :128].abs().max().item() == 0 and out_big[:, 512:].abs().max().item() == 0


# ---------------------------------------------------------------------------------------------
# RMSNorm / SwiGLU / RoPE
# ---------------------------------------------------------------------------------------------

@pytest.mark.parametrize("H", [256, 2048, 3072, 4096, 8192])
@pytest.mark.parametrize("rows", [1, 37, 4096])
def test_rmsnorm_fwd_bwd(rows, H):
    from neuronx_distributed_llama3_2_b200.ops.norm import fused_add_rms_norm, rms_norm, rms_norm_ref

    _lib()
    x = torch.randn(rows, H, device="cuda").bfloat16().requires_grad_()
    w = (1 + 0.1 * torch.randn(H, device="cuda")).bfloat16().requires_grad_()
    dy = torch.randn(rows, H, device="cuda").bfloat16()
    y = rms_norm(x, w, 1e-5)
    y.backward(dy)
    xr = x.detach().float().requires_grad_()
    wr = w.detach().float().requires_grad_()
    yr = rms_norm_ref(xr, wr, 1e-5)
    yr.backward(dy.float())
    assert _rel_err(y, yr) < 6e-3
    assert _rel_err(x.grad, xr.grad) < 8e-3
    assert _rel_err(w.grad, wr.grad) < 8e-3

    res = torch.randn(rows, H, device="cuda").bfloat16()
    y2, s2 = fused_add_rms_norm(x.detach(), res, w.detach(), 1e-5)
    s_ref = (x.detach().float() + res.float()).bfloat16()
    assert torch.equal(s2, s_ref)
    assert _rel_err(y2, rms_norm_ref(s_ref.float(), w.detach().float(), 1e-5)) < 6e-3


def test_fused_add_rmsnorm_backward():
    from neuronx_distributed_llama3_2_b200.ops.norm import fused_add_rms_norm, rms_norm_ref

    _lib()
    rows, H = 300, 2048
    x = torch.randn(rows, H, device="cuda").bfloat16().requires_grad_()
    r = torch.randn(rows, H, device="cuda").bfloat16().requires_grad_()
    w = torch.randn(H, device="cuda").bfloat16().requires_grad_()
    y, s = fused_add_rms_norm(x, r, w, 1e-5)
    gy, gs = torch.randn_like(y), torch.randn_like(s)
    (y.float() * gy.float()).sum().add((s.float() * gs.float()).sum()).backward()
    xf, rf, wf = (t.detach().float().requires_grad_() for t in (x, r, w))
    sf = xf + rf
    yf = rms_norm_ref(sf, wf, 1e-5)
    ((yf * gy.float()).sum() + (sf * gs.float()).sum()).backward()
    assert _rel_err(x.grad, xf.grad) < 1e-2
    assert _rel_err(r.grad, rf.grad) < 1e-2
    assert _rel_err(w.grad, wf.grad) < 1e-2


@pytest.mark.parametrize("I", [1024, 8192])
def test_swiglu(I):
    from neuronx_distributed_llama3_2_b200.ops.activation import swiglu, swiglu_ref

    _lib()
    gu = torch.randn(777, 2 * I, device="cuda").bfloat16().requires_grad_()
    dy = torch.randn(777, I, device="cuda").bfloat16()
    out = swiglu(gu)
    out.backward(dy)
    gr = gu.detach().float().requires_grad_()
    outr = swiglu_ref(gr)
    outr.backward(dy.float())
    assert _rel_err(out, outr) < 6e-3
    assert _rel_err(gu.grad, gr.grad) < 8e-3


@pytest.mark.parametrize("D", [64, 128])
def test_rope(D):
    from neuronx_distributed_llama3_2_b200.ops.rope import apply_rope, build_rope_cache, rope_ref

    _lib()
    S, B, heads, nrot = 128, 2, 6, 4
    cos, sin = build_rope_cache(S, D, 500000.0, {"rope_type": "llama3", "factor": 32.0}, device="cuda")
    x = torch.randn(S * B, heads, D, device="cuda").bfloat16().requires_grad_()
    y = apply_rope(x, cos, sin, nrot, pos_div=B, pos_mod=S)
    ref = rope_ref(x.detach().float(), cos, sin, nrot, B, S)
    assert _rel_err(y, ref) < 6e-3
    dy = torch.randn_like(y)
    y.backward(dy)
    ref_g = rope_ref(dy.float(), cos, sin, nrot, B, S, inverse=True)
    assert _rel_err(x.grad, ref_g) < 6e-3
    pos = torch.randint(0, S, (S * B,), device="cuda")
    y2 = apply_rope(x.detach(), cos, sin, nrot, positions=pos)
    assert _rel_err(y2, rope_ref(x.detach().float(), cos, sin, nrot, 1, 1, positions=pos)) < 6e-3


# ---------------------------------------------------------------------------------------------
# cross entropy, embedding
# ---------------------------------------------------------------------------------------------

def test_cross_entropy_kernels():
    from neuronx_distributed_llama3_2_b200.ops import cross_entropy as ce

    _lib()
    T, V, start = 513, 16032, 16032
    logits = (3 * torch.randn(T, V, device="cuda")).bfloat16()
    target = torch.randint(0, 3 * V, (T,), device="cuda")
    target[5] = -100
    stats = ce.ce_local_stats(logits, target, start)
    z = logits.float()
    m = z.max(-1).values
    assert torch.allclose(stats[:, 0], m)
    assert _rel_err(stats[:, 1], torch.exp(z - m[:, None]).sum(-1)) < 1e-4
    local = target - start
    ok = (local >= 0) & (local < V)
    tl = torch.where(ok, z.gather(1, local.clamp(0, V - 1)[:, None]).squeeze(1), torch.zeros_like(m))
    assert torch.allclose(stats[:, 2], tl)
    assert _rel_err(stats[:, 3], z.sum(-1)) < 1e-3
    lse = m + torch.log(stats[:, 1])
    go = torch.rand(T, device="cuda")
    for smoothing in (0.0, 0.1):
        grad = ce.ce_backward(logits, target, lse, go, start, smoothing, 3 * V)
        p = torch.exp(z - lse[:, None]) - smoothing / (3 * V)
        onehot = torch.zeros_like(p)
        onehot.scatter_(1, local.clamp(0, V - 1)[:, None], ok[:, None].float())
        ref = (p - (1 - smoothing) * onehot) * torch.where(target < 0, torch.zeros_like(go), go)[:, None]
        assert _rel_err(grad, ref) < 8e-3


def test_embedding_kernels():
    from neuronx_distributed_llama3_2_b200.ops.embedding import vocab_parallel_embedding

    _lib()
    V, H, B, S, start = 1000, 512, 3, 64, 500
    w = torch.randn(V, H, device="cuda").bfloat16().requires_grad_()
    ids = torch.randint(0, 2000, (B, S), device="cuda")
    for seq_first in (False, True):
        w.grad = None
        out = vocab_parallel_embedding(ids, w, start, seq_first_out=seq_first)
        local = ids - start
        mask = (local < 0) | (local >= V)
        ref = torch.nn.functional.embedding(local.masked_fill(mask, 0), w.detach()).masked_fill(mask[..., None], 0)
        if seq_first:
            ref = ref.transpose(0, 1)
        assert torch.equal(out, ref.contiguous())
        dy = torch.randn_like(out)
        out.backward(dy)
        wr = w.detach().float().requires_grad_()
        refo = torch.nn.functional.embedding(local.masked_fill(mask, 0), wr).masked_fill(mask[..., None], 0)
        if seq_first:
            refo = refo.transpose(0, 1)
        refo.backward(dy.float())
        assert _rel_err(w.grad, wr.grad) < 8e-3


# ---------------------------------------------------------------------------------------------
# optimizer kernels
# ---------------------------------------------------------------------------------------------

def test_optimizer_kernels():
    from neuronx_distributed_llama3_2_b200.ops import optim

    _lib()
    n = 3_000_017
    for dt in (torch.float32, torch.bfloat16):
        x = torch.randn(n, device="cuda").to(dt)
        out = torch.zeros(1, device="cuda")
        optim.sumsq_(x, out)
        assert abs(out.item() - x.float().pow(2).sum().item()) / n < 1e-3
        optim.sumsq_(x, out, accumulate=True)
        assert abs(out.item() - 2 * x.float().pow(2).sum().item()) / n < 2e-3
    coef = torch.empty(1, device="cuda")
    norm = torch.empty(1, device="cuda")
    tot = torch.tensor([25.0], device="cuda")
    optim.clip_coef_(tot, 1.0, coef, norm)
    assert abs(norm.item() - 5.0) < 1e-5 and abs(coef.item() - 1.0 / (5.0 + 1e-6)) < 1e-6
    optim.clip_coef_(torch.tensor([0.01], device="cuda"), 1.0, coef, norm)
    assert coef.item() == 1.0

    master = torch.randn(n, device="cuda")
    m = torch.zeros(n, device="cuda")
    v = torch.zeros(n, device="cuda")
    ref_p = master.clone().requires_grad_()
    opt = torch.optim.AdamW([ref_p], lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1)
    pout = torch.empty(n, device="cuda", dtype=torch.bfloat16)
    for step in range(1, 4):
        g = torch.randn(n, device="cuda").bfloat16()
        cc = torch.tensor([0.5], device="cuda")
        optim.adamw_step_(master, m, v, g, pout, 1e-2, 0.9, 0.95, 1e-8, 0.1, step, grad_scale=2.0, clip_coef=cc)
        ref_p.grad = g.float()  # 2.0 * 0.5 = 1
        opt.step()
        assert _rel_err(master, ref_p.detach()) < 1e-5
        assert torch.equal(pout, master.bfloat16())
    x = torch.randn(1001, device="cuda").bfloat16()
    x0 = x.clone()
    optim.scale_(x, 0.5, torch.tensor([4.0], device="cuda"))
    assert _rel_err(x, x0.float() * 2) < 1e-2


# ---------------------------------------------------------------------------------------------
# sampling
# ---------------------------------------------------------------------------------------------

@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("n,split_at", [(8192, 0), (8192, 8192), (100_000, 12_345), (3_000_000, 1_000_003)])
def test_zero1_reduce_shard_with_fused_sumsq(dt, n, split_at):
    """the ZeRO-1 pull-reduce with the sum of squares folded in, run with a one-rank "group" (the peer table
    holds the local buffer): the shard equals the fp32 cast of the gradient slice, the two sums of squares (before /
    from ``split_at`` on, any alignment) equal the fp64 sums and ACCUMULATE into their outputs, bitwise reproducible."""
    import ctypes as c

    lib = _lib()
    n = n // 8 * 8
    lo = 1024
    g = (torch.randn(lo + n + 64, device="cuda") * 3).to(dt)
    out = torch.empty(n, device="cuda", dtype=torch.float32)
    dtc = 1 if dt == torch.bfloat16 else 0
    blocks = int(lib.lib().nxd_zero1_reduce_blocks(c.c_longlong(n), c.c_int(dtc)))
    partials = torch.zeros(2 * blocks, device="cuda", dtype=torch.float32)
    sums = []
    for _ in range(2):
        s0 = torch.full((1,), 5.0, device="cuda")
        s1 = torch.full((1,), 7.0, device="cuda")
        ptrs = (c.c_void_p * 1)(g.data_ptr())
        rc = lib.lib().nxd_zero1_reduce_shard_sumsq(ptrs, c.c_int(dtc), c.c_longlong(lo), c.c_longlong(n), lib.ptr(out),
                                                    c.c_int(1), c.c_int(0), c.c_longlong(split_at), lib.ptr(partials),
                                                    lib.ptr(s0), lib.ptr(s1), lib.stream_ptr())
        lib.check(rc, "nxd_zero1_reduce_shard_sumsq")
        sums.append((float(s0), float(s1)))
    ref = g[lo:lo + n].float()
    assert torch.equal(out, ref)
    r0 = float(ref[:split_at].double().pow(2).sum()) + 5.0
    r1 = float(ref[split_at:].double().pow(2).sum()) + 7.0
    assert abs(sums[0][0] - r0) <= 1e-4 * max(1.0, r0) and abs(sums[0][1] - r1) <= 1e-4 * max(1.0, r1), (sums, r0, r1)
    assert sums[0] == sums[1]


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
def test_sampling(dt):
    from neuronx_distributed_llama3_2_b200.ops import sampling

    _lib()
    B, V = 9, 128256
    logits = torch.randn(B, V, device="cuda").to(dt)
    assert torch.equal(sampling.argmax(logits), torch.argmax(logits.float(), -1))
    u = torch.rand(B, device="cuda")
    for k, p in ((1, 1.0), (50, 1.0), (256, 0.9)):
        got = sampling.topk_sample(logits, k, p, 0.8, u)
        ref = sampling.topk_sample_ref(logits, k, p, 0.8, u)
        # bf16 ties / fp rounding at the CDF boundary may move a pick by one slot: require membership in top-k
        topk = torch.topk(logits.float(), k, dim=-1).indices
        assert all(int(got[b]) in topk[b].tolist() for b in range(B))
        assert (got == ref).float().mean().item() >= 0.75, (k, p, got, ref)


# ---------------------------------------------------------------------------------------------
# decode attention + KV cache
# ---------------------------------------------------------------------------------------------

@pytest.mark.parametrize("D", [64, 128])
def test_decode_attention_contiguous_and_paged(D):
    from neuronx_distributed_llama3_2_b200.ops import attention as A

    _lib()
    B, Hq, Hkv, Smax = 5, 8, 2, 4096
    ck = torch.randn(B, Hkv, Smax, D, device="cuda").bfloat16()
    cv = torch.randn(B, Hkv, Smax, D, device="cuda").bfloat16()
    q = torch.randn(B, Hq, D, device="cuda").bfloat16()
    lens = torch.tensor([1, 17, 1000, 4096, 2049], device="cuda")
    out = A.decode_attention(q, ck, cv, lens)
    ref = A.decode_attention_ref(q, ck, cv, lens, 1.0 / math.sqrt(D))
    assert _rel_err(out, ref) < 1e-2

    page = 128
    n_pages = B * Smax // page
    perm = torch.randperm(n_pages, device="cuda")
    table = perm.reshape(B, Smax // page).int()
    pk = torch.empty(n_pages, Hkv, page, D, device="cuda", dtype=torch.bfloat16)
    pv = torch.empty_like(pk)
    for b in range(B):
        for j in range(Smax // page):
            pk[table[b, j]] = ck[b, :, j * page:(j + 1) * page]
            pv[table[b, j]] = cv[b, :, j * page:(j + 1) * page]
    out_p = A.decode_attention(q, pk, pv, lens, block_table=table, page_size=page)
    assert _rel_err(out_p, ref) < 1e-2

    # KV write: T new tokens at position p of line seq_ids[b]
    nk = torch.randn(B, Hkv, 3, D, device="cuda").bfloat16()
    nv = torch.randn(B, Hkv, 3, D, device="cuda").bfloat16()
    pos = torch.tensor([0, 5, 100, 4093, 7], device="cuda")
    ck2, cv2 = ck.clone(), cv.clone()
    A.kv_cache_update(ck2, cv2, nk, nv, pos)
    for b in range(B):
        assert torch.equal(ck2[b, :, int(pos[b]):int(pos[b]) + 3], nk[b])
        assert torch.equal(cv2[b, :, int(pos[b]):int(pos[b]) + 3], nv[b])
    pk2, pv2 = pk.clone(), pv.clone()
    A.kv_cache_update(pk2, pv2, nk, nv, pos, block_table=table, page_size=page)
    for b in range(B):
        for t in range(3):
            p = int(pos[b]) + t
            assert torch.equal(pk2[table[b, p // page], :, p % page], nk[b, :, t])


@pytest.mark.parametrize("D", [64, 128])
@pytest.mark.parametrize("group", [1, 2, 3, 4, 8, 16, 32])
def test_decode_attention_gqa_groups(group, D):
    """Groups of 2 .. 16 query heads per KV head take the tensor-core kernel (heads = rows of the mma tile, rows past the
    group are zero), 1 and > 16 the FMA kernel; lengths around the 32-key tile and the split boundaries, and large score
    magnitudes (the running max must rescale)."""
    from neuronx_distributed_llama3_2_b200.ops import attention as A

    _lib()
    torch.manual_seed(group * 10 + D)
    Hkv, Smax = 2, 1024
    lens = torch.tensor([1, 31, 32, 33, 127, 129, 500, 1024, 1023], device="cuda")
    B = lens.numel()
    ck = torch.randn(B, Hkv, Smax, D, device="cuda").bfloat16()
    cv = torch.randn(B, Hkv, Smax, D, device="cuda").bfloat16()
    ck[:, :, 100:110] *= 4.0                       # a few dominant keys in the middle of the sequence
    cv[:, :, Smax - 8:] = float("nan")            # rows past every length < Smax must never be read into the sum
    lens_safe = torch.minimum(lens, torch.tensor(Smax - 8, device="cuda"))
    q = torch.randn(B, Hkv * group, D, device="cuda").bfloat16()
    out = A.decode_attention(q, ck, cv, lens_safe)
    ref = A.decode_attention_ref(q, ck, cv, lens_safe, 1.0 / math.sqrt(D))
    assert torch.isfinite(out.float()).all()
    assert _rel_err(out, ref) < 1e-2
    for b in range(B):
        assert _rel_err(out[b], ref[b]) < 2e-2, (b, int(lens_safe[b]))
    # continuous batching: row b reads cache line seq_ids[b] in place (lines repeated, as in speculation verify)
    lines = torch.tensor([3, 3, 0, 8, 1, 1, 1, 5, 2], device="cuda", dtype=torch.int32)
    out_l = A.decode_attention(q, ck, cv, lens_safe, seq_ids=lines)
    ref_l = A.decode_attention_ref(q, ck.index_select(0, lines.long()), cv.index_select(0, lines.long()), lens_safe,
                                   1.0 / math.sqrt(D))
    assert _rel_err(out_l, ref_l) < 1e-2


# ---------------------------------------------------------------------------------------------
# flash attention fwd / bwd
# ---------------------------------------------------------------------------------------------

@pytest.mark.parametrize("impl", ["tc", "mma"])
@pytest.mark.parametrize("D", [64, 128])
@pytest.mark.parametrize("S", [128, 1024, 1000, 2176])
def test_flash_attention(S, D, impl, monkeypatch):
    from neuronx_distributed_llama3_2_b200.ops import attention as A

    _lib()
    monkeypatch.setenv("NXD_ATTN_FWD", impl)
    monkeypatch.setenv("NXD_ATTN_BWD", impl)
    B, Hq, Hkv = 2, 6, 2
    q = torch.randn(B, Hq, S, D, device="cuda").bfloat16().requires_grad_()
    k = torch.randn(B, Hkv, S, D, device="cuda").bfloat16().requires_grad_()
    v = torch.randn(B, Hkv, S, D, device="cuda").bfloat16().requires_grad_()
    do = torch.randn(B, Hq, S, D, device="cuda").bfloat16()
    assert A._backend(q) == "nxd", "the hand-written flash attention must be the GPU path"
    out = A.flash_attn_func(q, k, v, causal=True)
    out.backward(do)
    qf, kf, vf = (t.detach().float().requires_grad_() for t in (q, k, v))
    ref = A.attention_ref(qf, kf, vf, True, None)
    ref.backward(do.float())
    assert _rel_err(out, ref) < 1e-2
    assert _rel_err(q.grad, qf.grad) < 2e-2
    assert _rel_err(k.grad, kf.grad) < 2e-2
    assert _rel_err(v.grad, vf.grad) < 2e-2


def test_flash_attention_strided_layout_noncausal():
    """[S, B, heads, D] storage viewed as [B, H, S, D] (how the SP transformer block holds q/k/v)."""
    from neuronx_distributed_llama3_2_b200.ops import attention as A

    _lib()
    S, B, Hq, Hkv, D = 512, 2, 4, 1, 128
    qkv = torch.randn(S, B, Hq + 2 * Hkv, D, device="cuda").bfloat16()
    q = qkv[:, :, :Hq].permute(1, 2, 0, 3)
    k = qkv[:, :, Hq:Hq + Hkv].permute(1, 2, 0, 3)
    v = qkv[:, :, Hq + Hkv:].permute(1, 2, 0, 3)
    for impl in ("tc", "mma"):
        import os
        os.environ["NXD_ATTN_FWD"] = impl
        for causal in (True, False):
            out = A.flash_attn_func(q, k, v, causal=causal)
            ref = A.attention_ref(q.float(), k.float(), v.float(), causal, None)
            assert _rel_err(out, ref) < 1e-2, (impl, causal, _rel_err(out, ref))
    os.environ.pop("NXD_ATTN_FWD", None)


@pytest.mark.parametrize("split", ["1", "0"])
@pytest.mark.parametrize("D", [128, 64])
@pytest.mark.parametrize("causal", [True, False])
def test_flash_attention_bwd_tc_gqa_cross_lengths(causal, D, split, monkeypatch):
    """tcgen05 backward: GQA group of 3, Sq != Sk, ragged lengths, strided [S, B, H, D] storage; with and without
    several CTAs sharing a K/V block (partial dK / dV reduce-added into fp32 accumulators)."""
    from neuronx_distributed_llama3_2_b200.ops import attention as A

    _lib()
    monkeypatch.setenv("NXD_ATTN_BWD_SPLIT", split)
    B, Hq, Hkv, Sq, Sk = 2, 6, 2, 1224, 1352
    qs = torch.randn(Sq, B, Hq, D, device="cuda").bfloat16().requires_grad_()
    ks = torch.randn(Sk, B, Hkv, D, device="cuda").bfloat16().requires_grad_()
    vs = torch.randn(Sk, B, Hkv, D, device="cuda").bfloat16().requires_grad_()
    q, k, v = (t.permute(1, 2, 0, 3) for t in (qs, ks, vs))
    do = torch.randn(B, Hq, Sq, D, device="cuda").bfloat16()
    out = A.flash_attn_func(q, k, v, causal=causal)
    out.backward(do)
    qf, kf, vf = (t.detach().float().requires_grad_() for t in (q, k, v))
    ref = A.attention_ref(qf, kf, vf, causal, None)
    ref.backward(do.float())
    assert _rel_err(out, ref) < 1e-2
    assert _rel_err(qs.grad.permute(1, 2, 0, 3), qf.grad) < 2e-2
    assert _rel_err(ks.grad.permute(1, 2, 0, 3), kf.grad) < 2e-2
    assert _rel_err(vs.grad.permute(1, 2, 0, 3), vf.grad) < 2e-2


def test_flash_attention_large_magnitude_rows():
    """Scores with a wide dynamic range exercise the lazy O rescale of the tcgen05 kernel."""
    from neuronx_distributed_llama3_2_b200.ops import attention as A

    _lib()
    B, Hq, Hkv, S, D = 1, 2, 1, 1024, 128
    q = (4 * torch.randn(B, Hq, S, D, device="cuda")).bfloat16()
    k = (4 * torch.randn(B, Hkv, S, D, device="cuda")).bfloat16()
    k[:, :, 700:] *= 3.0     # late keys dominate -> the running max jumps late in the loop
    v = torch.randn(B, Hkv, S, D, device="cuda").bfloat16()
    out = A.flash_attn_func(q, k, v, causal=False)
    ref = A.attention_ref(q.float(), k.float(), v.float(), False, None)
    assert _rel_err(out, ref) < 1.5e-2


@pytest.mark.parametrize("B,K,N", [(1, 2048, 512), (4, 3072, 1024), (5, 1024, 3080), (8, 4096, 256), (16, 2048, 1536), (9, 4096, 264),
                                   (13, 1032, 2048)])
def test_skinny_gemm(B, K, N):
    from neuronx_distributed_llama3_2_b200.ops.fused_comm import skinny_gemm

    x = torch.randn(B, K, device="cuda").bfloat16()
    w = (torch.randn(N, K, device="cuda") * 0.05).bfloat16()
    res = torch.randn(B, N, device="cuda").bfloat16()
    out = skinny_gemm(x, w)
    ref = x.float() @ w.float().t()
    assert out is not None
    assert ((out.float() - ref).norm() / ref.norm()).item() < 4e-3
    out = skinny_gemm(x, w, residual=res)
    ref = ref + res.float()
    assert ((out.float() - ref).norm() / ref.norm()).item() < 4e-3


@pytest.mark.parametrize("splits", [0, 1, 2, 4, 8])
@pytest.mark.parametrize("B,K,N", [(16, 2048, 3072), (9, 2048, 2048), (32, 8192, 2048), (33, 1032, 1000), (64, 2048, 16384),
                                   (100, 3072, 520), (128, 4096, 264), (1, 512, 128), (17, 576, 2048)])
def test_gemm_smallm(B, K, N, splits):
    """Weight-streaming tcgen05 GEMM (weight rows = UMMA M) incl. the cluster split-K reduction, ragged N / K, a K
    slice count that leaves the last cluster rank without k blocks (K = 576 -> 9 blocks over 4 / 8 ranks), strided x."""
    from neuronx_distributed_llama3_2_b200.ops.gemm import gemm_smallm

    xw = torch.randn(B, K + 64, device="cuda").bfloat16()
    x = xw[:, :K]                                    # row pitch K + 64
    w = (torch.randn(N, K, device="cuda") * 0.05).bfloat16()
    res = torch.randn(B, N, device="cuda").bfloat16()
    ref = x.float() @ w.float().t()
    out = gemm_smallm(x, w, splits=splits)
    assert out is not None
    assert ((out.float() - ref).norm() / ref.norm()).item() < 4e-3
    assert (out.float() - ref).abs().max().item() < 0.05 * ref.abs().max().item() + 0.05
    out = gemm_smallm(x, w, residual=res, splits=splits)
    ref = ref + res.float()
    assert ((out.float() - ref).norm() / ref.norm()).item() < 4e-3
    # bitwise reproducible (rank-ordered reduction, no atomics)
    assert torch.equal(out, gemm_smallm(x, w, residual=res, splits=splits))


@pytest.mark.parametrize("C", [384, 200, 37])
@pytest.mark.parametrize("layout", ["nn", "nt", "tn"])
def test_grouped_expert_gemm(layout, C):
    """MoE expert GEMMs (ONE grouped tcgen05 launch for all experts; ragged capacities overhang into out-of-bounds rows of
    the 3-D tensor maps) against torch.bmm in fp32, forward + autograd function."""
    from neuronx_distributed_llama3_2_b200.modules.moe.moe_parallel_layers import (
        ExpertFusedLinearWithAsyncCommunication, _grouped_mm)

    E, C, A, Bd = 4, C, 256, 512
    x = torch.randn(E, C, A, device="cuda").bfloat16()
    w = (torch.randn(E, A, Bd, device="cuda") * 0.05).bfloat16()
    if layout == "nn":
        got, ref = _grouped_mm(x, w, "nn"), torch.bmm(x.float(), w.float())
    elif layout == "nt":
        g = torch.randn(E, C, Bd, device="cuda").bfloat16()
        got, ref = _grouped_mm(g, w, "nt"), torch.bmm(g.float(), w.float().transpose(1, 2))
    else:
        g = torch.randn(E, C, Bd, device="cuda").bfloat16()
        got, ref = _grouped_mm(x, g, "tn"), torch.bmm(x.float().transpose(1, 2), g.float())
    assert _rel_err(got, ref) < 1e-2
    if layout == "nn":
        xr, wr = x.clone().requires_grad_(), w.clone().requires_grad_()
        y = ExpertFusedLinearWithAsyncCommunication.apply(xr, wr, False)
        dy = torch.randn_like(y)
        y.backward(dy)
        xf, wf = x.float().requires_grad_(), w.float().requires_grad_()
        torch.bmm(xf, wf).backward(dy.float())
        assert _rel_err(xr.grad, xf.grad) < 1e-2 and _rel_err(wr.grad, wf.grad) < 1e-2


@pytest.mark.parametrize("B,K,N", [(1, 2048, 3072), (2, 1024, 520), (4, 4096, 2048), (7, 2048, 1000), (8, 8192, 3072)])
@pytest.mark.parametrize("per_channel", [False, True])
def test_skinny_gemm_i8(B, K, N, per_channel):
    """int8 weight-only token-generation GEMM against the fp32 product with the dequantised weight."""
    from neuronx_distributed_llama3_2_b200.ops.fused_comm import skinny_gemm_i8

    x = torch.randn(B, K, device="cuda").bfloat16()
    w = torch.randint(-128, 128, (N, K), device="cuda", dtype=torch.int8)
    scale = (torch.rand(N if per_channel else 1, device="cuda") * 0.01 + 0.001).float()
    res = torch.randn(B, N, device="cuda").bfloat16()
    ref = (x.float() @ w.float().t()) * scale.reshape(1, -1)
    out = skinny_gemm_i8(x, w, scale)
    assert out is not None
    assert ((out.float() - ref).norm() / ref.norm()).item() < 4e-3
    out = skinny_gemm_i8(x, w, scale, residual=res)
    ref = ref + res.float()
    assert ((out.float() - ref).norm() / ref.norm()).item() < 4e-3


@pytest.mark.parametrize("B,D,paged", [(1, 64, False), (3, 128, False), (8, 64, True)])
def test_decode_fused_qkv_rope_cache(B, D, paged):
    """RMSNorm + fused QKV GEMV + RoPE + KV-cache write in one launch against the unfused fp32 math."""
    from neuronx_distributed_llama3_2_b200.ops import decode_fused, rope as rope_ops

    torch.manual_seed(B * 7 + D)
    H, nq, nkv, S = 1024, 8, 2, 256
    x = torch.randn(B, H, device="cuda").bfloat16()
    nw = (1 + 0.1 * torch.randn(H, device="cuda")).bfloat16()
    w = (torch.randn((nq + 2 * nkv) * D, H, device="cuda") * 0.03).bfloat16()
    cos, sin = rope_ops.build_rope_cache(S, D, 500000.0, None, device="cuda")
    pos = torch.randint(0, S, (B,), device="cuda", dtype=torch.int32)
    page = 16
    if paged:
        n_pages = B * S // page
        ck = torch.zeros(n_pages, nkv, page, D, device="cuda", dtype=torch.bfloat16)
        cv = torch.zeros_like(ck)
        table = torch.randperm(n_pages, device="cuda").reshape(B, S // page).to(torch.int32)
    else:
        ck = torch.zeros(B, nkv, S, D, device="cuda", dtype=torch.bfloat16)
        cv = torch.zeros_like(ck)
        table = None
    q = decode_fused.qkv_rope_cache(x, nw, 1e-5, w, cos, sin, pos, ck, cv, nq, nkv, D, block_table=table,
                                    page_size=page if paged else 0)
    xf = x.float()
    xn = (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + 1e-5) * nw.float()).bfloat16().float()
    qkv = (xn @ w.float().t()).reshape(B, nq + 2 * nkv, D)
    c, s_ = cos[pos.long()], sin[pos.long()]                       # [B, D/2]
    a, b_ = qkv[:, : nq + nkv, : D // 2], qkv[:, : nq + nkv, D // 2:]
    rot = torch.cat([a * c[:, None] - b_ * s_[:, None], b_ * c[:, None] + a * s_[:, None]], dim=-1)
    assert (q.float() - rot[:, :nq]).abs().max() < 3e-2
    for b in range(B):
        p = int(pos[b])
        if paged:
            krow = ck[int(table[b, p // page]), :, p % page]
            vrow = cv[int(table[b, p // page]), :, p % page]
        else:
            krow, vrow = ck[b, :, p], cv[b, :, p]
        assert (krow.float() - rot[b, nq:]).abs().max() < 3e-2
        assert (vrow.float() - qkv[b, nq + nkv:]).abs().max() < 3e-2
    assert int((ck != 0).any(-1).sum()) == B * nkv        # exactly one row per (sequence, kv head) was written


@pytest.mark.parametrize("splits", [0, 1, 4])
@pytest.mark.parametrize("B,D,nq,nkv,paged", [(16, 64, 8, 2, False), (9, 128, 6, 2, False), (33, 64, 7, 1, True),
                                              (64, 128, 4, 4, False), (100, 64, 8, 2, True)])
def test_smallm_qkv_rope_cache(B, D, nq, nkv, paged, splits):
    """Batched (9 .. 128 rows) fused QKV tensor-core GEMM + RoPE + KV-cache write against the unfused fp32 math; D = 64
    with an odd head count leaves the last 128-row weight slab half empty."""
    from neuronx_distributed_llama3_2_b200.ops import decode_fused, rope as rope_ops

    torch.manual_seed(B * 7 + D)
    H, S = 1024, 256
    xn = torch.randn(B, H, device="cuda").bfloat16()
    w = (torch.randn((nq + 2 * nkv) * D, H, device="cuda") * 0.03).bfloat16()
    cos, sin = rope_ops.build_rope_cache(S, D, 500000.0, None, device="cuda")
    pos = torch.randint(0, S, (B,), device="cuda", dtype=torch.int32)
    page = 16
    if paged:
        n_pages = B * S // page
        ck = torch.zeros(n_pages, nkv, page, D, device="cuda", dtype=torch.bfloat16)
        cv = torch.zeros_like(ck)
        table = torch.randperm(n_pages, device="cuda").reshape(B, S // page).to(torch.int32)
    else:
        ck = torch.zeros(B + 3, nkv, S, D, device="cuda", dtype=torch.bfloat16)     # more cache lines than rows
        cv = torch.zeros_like(ck)
        table = None
    # continuous batching: row b writes cache line seq_ids[b] (a permutation that is not the identity)
    lines = torch.randperm(B if paged else B + 3, device="cuda")[:B].to(torch.int32)
    q = decode_fused.qkv_rope_cache_batched(xn, w, cos, sin, pos, ck, cv, nq, nkv, D, seq_ids=lines, block_table=table,
                                            page_size=page if paged else 0, splits=splits)
    qkv = (xn.float() @ w.float().t()).reshape(B, nq + 2 * nkv, D)
    c, s_ = cos[pos.long()], sin[pos.long()]                       # [B, D/2]
    a, b_ = qkv[:, : nq + nkv, : D // 2], qkv[:, : nq + nkv, D // 2:]
    rot = torch.cat([a * c[:, None] - b_ * s_[:, None], b_ * c[:, None] + a * s_[:, None]], dim=-1)
    assert (q.float() - rot[:, :nq]).abs().max() < 3e-2
    for b in range(B):
        p = int(pos[b])
        line = int(lines[b])
        if paged:
            krow = ck[int(table[line, p // page]), :, p % page]
            vrow = cv[int(table[line, p // page]), :, p % page]
        else:
            krow, vrow = ck[line, :, p], cv[line, :, p]
        assert (krow.float() - rot[b, nq:]).abs().max() < 3e-2
        assert (vrow.float() - qkv[b, nq + nkv:]).abs().max() < 3e-2
    assert int((ck != 0).any(-1).sum()) == B * nkv        # exactly one row per (sequence, kv head) was written


@pytest.mark.parametrize("splits", [0, 1, 2, 8])
@pytest.mark.parametrize("B,H,I", [(16, 2048, 8192), (9, 1024, 1000), (40, 3072, 2048), (64, 1024, 520)])
def test_smallm_gate_up_swiglu(B, H, I, splits):
    """Batched gate/up tensor-core GEMM (two accumulators per CTA) + SwiGLU against fp32; I = 1000 / 520 end inside a
    128-row slab (the gate slab then overlaps the first up rows, which must not leak into the output)."""
    from neuronx_distributed_llama3_2_b200.ops import decode_fused

    torch.manual_seed(B + I)
    xn = torch.randn(B, H, device="cuda").bfloat16()
    w = (torch.randn(2 * I, H, device="cuda") * 0.03).bfloat16()
    out = decode_fused.gate_up_swiglu_batched(xn, w, splits=splits)
    gu = (xn.float() @ w.float().t()).bfloat16().float()
    ref = torch.nn.functional.silu(gu[:, :I]) * gu[:, I:]
    assert ((out.float() - ref).norm() / ref.norm()).item() < 6e-3
    assert (out.float() - ref).abs().max().item() < 0.05 * ref.abs().max().item() + 0.02


@pytest.mark.parametrize("B", [1, 5, 8])
def test_decode_fused_gate_up_swiglu(B):
    from neuronx_distributed_llama3_2_b200.ops import decode_fused

    torch.manual_seed(B)
    H, I = 2048, 1536
    x = torch.randn(B, H, device="cuda").bfloat16()
    nw = (1 + 0.1 * torch.randn(H, device="cuda")).bfloat16()
    w = (torch.randn(2 * I, H, device="cuda") * 0.03).bfloat16()
    out = decode_fused.gate_up_swiglu(x, nw, 1e-5, w)
    xf = x.float()
    xn = (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + 1e-5) * nw.float()).bfloat16().float()
    gu = xn @ w.float().t()
    ref = torch.nn.functional.silu(gu[:, :I]) * gu[:, I:]
    assert (out.float() - ref).abs().max() < 3e-2 * max(1.0, float(ref.abs().max()))
    plain = decode_fused.gate_up_swiglu(x, None, 0.0, w)           # no norm: x is used as is
    gu2 = xf @ w.float().t()
    ref2 = torch.nn.functional.silu(gu2[:, :I]) * gu2[:, I:]
    assert (plain.float() - ref2).abs().max() < 3e-2 * max(1.0, float(ref2.abs().max()))


@pytest.mark.parametrize("D,causal", [(128, True), (64, True), (128, False)])
def test_flash_attention_dropout_matches_oracle_mask(D, causal):
    """Attention dropout in the tcgen05 kernels (reference ``kernels/flash_attn.py:122-127,151-160``): forward and all three
    gradients against fp32 math that applies the SAME keep mask (restated from the kernel's hash), plus the keep rate and
    seed dependence."""
    from neuronx_distributed_llama3_2_b200.ops import attention as A
    from neuronx_distributed_llama3_2_b200.ops.flash_attn import dropout_keep_mask

    torch.manual_seed(D)
    B, Hq, Hkv, S, p_drop, seed = 2, 4, 2, 320, 0.2, 987654321
    q = (torch.randn(B, Hq, S, D, device="cuda") * 0.5).bfloat16().requires_grad_()
    k = (torch.randn(B, Hkv, S, D, device="cuda") * 0.5).bfloat16().requires_grad_()
    v = torch.randn(B, Hkv, S, D, device="cuda").bfloat16().requires_grad_()
    do = torch.randn(B, Hq, S, D, device="cuda").bfloat16()
    o = A.flash_attn_func(q, k, v, dropout_p=p_drop, causal=causal, seed=seed)
    o.backward(do)
    got = (o.detach().float(), q.grad.float(), k.grad.float(), v.grad.float())
    qf, kf, vf = (t.detach().float().requires_grad_() for t in (q, k, v))
    ref = A.attention_ref(qf, kf, vf, causal, None, p_drop, seed).float()
    ref.backward(do.float())
    want = (ref.detach(), qf.grad, kf.grad, vf.grad)
    for name, a, b in zip(("o", "dq", "dk", "dv"), got, want):
        rel = float((a - b).norm() / (b.norm() + 1e-12))
        assert rel < 2e-2, (name, rel)
    keep = dropout_keep_mask(seed, B, Hq, S, S, p_drop, device="cuda")
    assert abs(float(keep.float().mean()) - (1 - p_drop)) < 5e-3
    o2 = A.flash_attn_func(q.detach(), k.detach(), v.detach(), dropout_p=p_drop, causal=causal, seed=seed + 1)
    assert float((o2.float() - o.detach().float()).abs().max()) > 1e-2      # another seed, another mask
    o3 = A.flash_attn_func(q.detach(), k.detach(), v.detach(), dropout_p=p_drop, causal=causal, seed=seed)
    assert torch.equal(o3, o.detach())



def test_moe_capacity_path_permute_kernels_match_indexing_path(monkeypatch):
    """forward_capacity_factor through the gather / combine kernels + grouped GEMM (with token dropping, ragged capacity)
    against the same module on the torch indexing path in fp32: output, input gradient, expert and router gradients."""
    from neuronx_distributed_llama3_2_b200.modules.moe import ExpertMLPs, RouterTopK
    from neuronx_distributed_llama3_2_b200.ops import moe_permute

    torch.manual_seed(3)
    T, H, I, E, k = 300, 256, 512, 4, 2
    def build(dtype):
        torch.manual_seed(11)
        router = RouterTopK(E, k, H, dtype=dtype, device=torch.device("cuda"))
        experts = ExpertMLPs(E, k, H, I, "silu", True, 1.0, normalize_top_k_affinities=True, dtype=dtype,
                             device=torch.device("cuda"))
        return router, experts
    r16, e16 = build(torch.bfloat16)
    r32, e32 = build(torch.float32)
    with torch.no_grad():
        for a, b in zip(list(r32.parameters()) + list(e32.parameters()), list(r16.parameters()) + list(e16.parameters())):
            a.copy_(b.float())
    x = torch.randn(T, H, device="cuda")
    x16 = x.bfloat16().requires_grad_()
    x32 = x.bfloat16().float().requires_grad_()
    e16.train(); e32.train()
    calls = []
    real = moe_permute._gather
    monkeypatch.setattr(moe_permute, "_gather", lambda *a: (calls.append(1), real(*a))[1])
    _, aff16, idx = r16(x16)
    out16 = e16(x16, aff16, idx, T)
    assert calls, "the permute kernels did not run"
    _, aff32, _ = r32(x32)
    out32 = e32(x32, aff32, idx, T)
    dy = torch.randn(T, H, device="cuda")
    out16.backward(dy.bfloat16())
    out32.backward(dy.bfloat16().float())
    assert _rel_err(out16, out32) < 2e-2
    assert _rel_err(x16.grad, x32.grad) < 3e-2
    for (n, a), (_, b) in zip(list(e16.named_parameters()) + list(r16.named_parameters()),
                              list(e32.named_parameters()) + list(r32.named_parameters())):
        assert _rel_err(a.grad, b.grad) < 4e-2, n
